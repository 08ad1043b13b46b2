// twc.cuh -- thread / warp / CTA cooperative forward advance for small frontiers.
//
// Replaces the reference's edge_map_forward (Merrill-style "TWC") expansion
// (gunrock/oprtr/edge_map_forward/cta.cuh:180-541,707-841, kernel.cuh:274-316:
// texture fetches of row offsets, SOA raking scans, 48 KB smem budget,
// thresholds 32 / 512) with one sm_100a kernel and no scratch memory:
//
//   * a CTA dequeues THREADS frontier items; each thread loads its item's row
//     range (the two offsets sit in one 8/16-byte span);
//   * lists >= CTA_THRESHOLD are gathered by the whole CTA, one list at a time
//     (owner elected through shared memory), 128-byte coalesced per warp;
//   * lists >= WARP_THRESHOLD are gathered by the owning warp (owner elected by
//     ballot, range broadcast by shuffle);
//   * the remaining short lists are packed with a CTA-wide exclusive scan of
//     their lengths into a shared-memory gather table so that consecutive
//     threads read consecutive table slots (no idle lanes on 1-3 edge lists).
//   Winners are enqueued with one ballot + one global atomic per warp step
//   (frontiers routed here are small, so that atomic traffic is negligible and no
//   CTA-wide staging barrier is needed inside the divergent warp phase).
//
// One launch, no host round trip: this is the path for the many short levels
// of high-diameter graphs and for the first / last levels of an R-MAT search.
// A single huge list (> ~64K entries) is better served by the LB mode, which
// spreads it over all SMs; the dispatcher (kernel.cuh) chooses.
#pragma once
#include "common.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {
namespace twc {

template <typename Policy, bool NEED_SRC>
struct Storage {
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    SizeT      gather_pos[Policy::TILE_EDGES];                  // fine-grained gather table: CSR position ...
    VertexId   gather_src[NEED_SRC ? Policy::TILE_EDGES : 1];   // ... source vertex ...
    int        gather_in[Policy::TILE_EDGES];                   // ... and thread (frontier slot) of each entry
    unsigned   warp_tot[Policy::THREADS / 32 + 1];
    unsigned long long red[Policy::THREADS / 32 + 1];
    int        owner;
    long long  cta_begin, cta_end, cta_in;
    VertexId   cta_src;
};

// Expands frontier tiles cta, cta + num_ctas, ... (THREADS items each).  Callable from any kernel
// whose CTAs have Policy::THREADS threads.
template <typename Policy, typename Problem, typename Functor>
__device__ __forceinline__ void ExpandItems(
    const typename Problem::DataSlice &ds,
    const Args<typename Policy::VertexId, typename Policy::SizeT> &args,
    Storage<Policy, Problem::MARK_PREDECESSORS> &sm,
    unsigned long long in_len, LevelCounters *slot_out, unsigned cta, unsigned num_ctas)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    constexpr int THREADS = Policy::THREADS;
    constexpr int GATHER  = Policy::TILE_EDGES;          // slots of the fine-grained gather table
    constexpr int ROUNDS  = GATHER / THREADS;
    constexpr bool NEED_SRC = Problem::MARK_PREDECESSORS;

    unsigned long long my_edges = 0, my_out_degree = 0;
    const int lane = util::LaneId();
    const bool compact = args.output == OUT_COMPACT;

    // one edge: functor pair + staging (warp-convergent; `live` masks idle lanes)
    auto visit = [&](bool live, long long pos, VertexId src, long long in_idx) {
        bool ok = false;
        VertexId nbr = -1;
        if (live) {
            nbr = util::LoadStream(args.g.col_indices + pos);
            ++my_edges;
            bool culled = false;
            if constexpr (Policy::PRE_CULL) culled = util::BitTest(ds.d_visited_mask, ds.BitPos(nbr));
            if (!culled) {
                VertexId s_id = NEED_SRC ? src : (VertexId)args.iteration;
                ok = VisitEdge<Policy, Problem, Functor>(&ds, s_id, nbr, (SizeT)pos, (SizeT)in_idx);
            }
        }
        if (compact) {
            const unsigned m = __ballot_sync(util::kFullMask, ok);
            if (m) {
                const int leader = __ffs(m) - 1;
                unsigned long long b = 0;
                if (lane == leader) b = atomicAdd(&slot_out->out_len, (unsigned long long)__popc(m));
                b = __shfl_sync(util::kFullMask, b, leader);
                if (ok) {
                    const unsigned long long at = b + __popc(m & ((1u << lane) - 1));
                    if (at < args.out_cap) args.out_queue[at] = OutputIsEdges(args.type) ? (VertexId)pos : nbr;
                    else slot_out->overflow = 1;
                    if constexpr (Policy::TRACK_OUT_DEGREE) {
                        if (args.g.part.Mine(nbr)) {
                            long long row = args.g.part.Local(nbr);
                            my_out_degree += (unsigned long long)(args.g.row_offsets[row + 1] - args.g.row_offsets[row]);
                        }
                    }
                }
            }
        }
    };

    const unsigned long long num_tiles = (in_len + THREADS - 1) / THREADS;
    for (unsigned long long tile = cta; tile < num_tiles; tile += num_ctas) {
        const unsigned long long i = tile * THREADS + threadIdx.x;
        VertexId v = -1;
        long long begin = 0, end = 0;
        if (i < in_len) {
            v = args.in_queue ? util::LoadCG(args.in_queue + i) : (VertexId)args.g.part.Global((long long)i);
            if (InputIsEdges(args.type) && args.in_queue)       // an edge id stands for the vertex it points to
                v = (v >= 0 && (long long)v < args.g.edges) ? args.g.col_indices[v] : (VertexId)-1;
            if (v >= 0 && (long long)v < args.g.nodes) {
                long long row = args.g.part.Local(v);
                begin = (long long)args.g.row_offsets[row];
                end   = (long long)args.g.row_offsets[row + 1];
            }
        }
        long long len = end - begin;

        // ---- CTA-cooperative: one long list at a time -------------------------------
        for (;;) {
            __syncthreads();
            if (threadIdx.x == 0) sm.owner = -1;
            __syncthreads();
            if (len >= Policy::CTA_THRESHOLD) sm.owner = threadIdx.x;
            __syncthreads();
            const int owner = sm.owner;
            if (owner < 0) break;
            if (owner == (int)threadIdx.x) {
                sm.cta_begin = begin; sm.cta_end = end; sm.cta_src = v; sm.cta_in = (long long)i;
                len = 0;
            }
            __syncthreads();
            const long long cb = sm.cta_begin, ce = sm.cta_end;
            const VertexId  cs = sm.cta_src;
            const long long ci = sm.cta_in;
            for (long long p = cb; p < ce; p += THREADS)       // uniform trip count
                visit(p + threadIdx.x < ce, p + threadIdx.x, cs, ci);
        }

        // ---- warp-cooperative: medium lists ---------------------------------------------
        {
            unsigned m = __ballot_sync(util::kFullMask, len >= Policy::WARP_THRESHOLD);
            while (m) {
                const int leader = __ffs(m) - 1;
                m &= m - 1;
                const long long wb = __shfl_sync(util::kFullMask, begin, leader);
                const long long we = __shfl_sync(util::kFullMask, end, leader);
                const VertexId  ws = __shfl_sync(util::kFullMask, v, leader);
                const long long wi = (long long)(tile * THREADS + (threadIdx.x & ~31) + leader);
                if (lane == leader) len = 0;
                for (long long p = wb; p < we; p += 32) visit(p + lane < we, p + lane, ws, wi);
            }
        }

        // ---- thread-level: short lists packed through the gather table --------------------
        {
            unsigned total;
            unsigned excl = util::BlockExclusiveSum<THREADS>((unsigned)len, sm.warp_tot, total);
            for (unsigned round_base = 0; round_base < total; round_base += GATHER) {
                for (unsigned k = 0; k < (unsigned)len; ++k) {
                    unsigned slot = excl + k - round_base;         // wraps to huge if below the window
                    if (slot < (unsigned)GATHER) {
                        sm.gather_pos[slot] = (SizeT)(begin + k);
                        sm.gather_in[slot]  = threadIdx.x;
                        if constexpr (NEED_SRC) sm.gather_src[slot] = v;
                    }
                }
                __syncthreads();
                const unsigned n_round = (total - round_base < (unsigned)GATHER) ? (total - round_base) : (unsigned)GATHER;
#pragma unroll
                for (int r = 0; r < ROUNDS; ++r) {
                    const unsigned slot = r * THREADS + threadIdx.x;
                    const bool live = slot < n_round;
                    long long pos = live ? (long long)sm.gather_pos[slot] : 0;
                    VertexId  src = (NEED_SRC && live) ? sm.gather_src[slot] : (VertexId)0;
                    long long in_idx = live ? (long long)(tile * THREADS + sm.gather_in[slot]) : 0;
                    visit(live, pos, src, in_idx);
                }
                __syncthreads();                                // table is rewritten next round
            }
        }
    }
    BlockAccumulate<THREADS>(my_edges, sm.red, &slot_out->edges);
    if constexpr (Policy::TRACK_OUT_DEGREE) BlockAccumulate<THREADS>(my_out_degree, sm.red, &slot_out->out_degree);
}

template <typename Policy, typename Problem, typename Functor>
__global__ void __launch_bounds__(Policy::THREADS, Policy::MIN_CTAS)
Kernel(const __grid_constant__ typename Problem::DataSlice ds,
       const __grid_constant__ Args<typename Policy::VertexId, typename Policy::SizeT> args)
{
    __shared__ Storage<Policy, Problem::MARK_PREDECESSORS> s_twc;
    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (blockIdx.x == 0 && threadIdx.x == 0) ClearFutureSlot(args.slots, args.queue_index);
    ExpandItems<Policy, Problem, Functor>(ds, args, s_twc, in_len, slot_out, blockIdx.x, gridDim.x);
}

}  // namespace twc
}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
