// pull.cuh -- pull-direction ("bottom-up", backward) advance for direction-optimising BFS.
//
// Replaces the reference's edge_map_backward / edge_map_partitioned_backward
// (gunrock/oprtr/edge_map_backward/cta.cuh:220-341,652-773;
// oprtr/edge_map_partitioned_backward/kernel.cuh:377-526), which keep the
// frontier as `bool` BYTES (dobfs_problem.cuh:63-64), materialise an
// unvisited-vertex queue from an iota array with two extra filter passes
// (dobfs_enactor.cuh:422-462) and have no cross-thread early exit.
//
// Here the state is three 1-bit-per-vertex maps (n/8 bytes each; 4 MiB at scale
// 25, resident in the 126 MB L2):
//     visited        read+written, only the words of the rows this device owns
//     frontier_in    read-only (probed through the non-coherent path)
//     frontier_out   written whole words, no atomics, no prior memset
// and the kernel SWEEPS the owned rows: one warp per 32-vertex word, one lane
// per vertex.  A word whose vertices are all visited (or edgeless: `candidates`)
// costs one 4-byte load.  An unvisited lane walks its in-neighbour list with
// early exit; lists longer than SERIAL_LIMIT are finished by the whole warp with
// 128-byte coalesced reads and a ballot early exit, so a straggler list cannot
// serialise a warp.  Newly found vertices are published as one word store.
//
// Algorithmic HBM bytes per call: n/8 (visited) [+ n/8 candidates] + n/8
// (frontier_out) for the sweep, 2*sizeof(SizeT) per unvisited candidate, 4 + 1
// per inspected neighbour entry (column id + frontier bit probe), 4 per found
// vertex (label).
#pragma once
#include "common.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {
namespace pull {

template <typename VertexId, typename SizeT>
struct PullArgs {
    graph::CsrView<VertexId, SizeT> g;      // INVERSE graph (CSC); == CSR when symmetric
    const unsigned *frontier_in;            // global bitmap, owner-major bit positions
    unsigned       *frontier_out;           // global bitmap; this device writes its own words
    unsigned       *visited;                // global bitmap; this device updates its own words
    const unsigned *candidates;             // optional, local words: 1 = vertex has in-edges
    const SizeT    *fwd_row_offsets;        // forward (out-edge) offsets for the degree sum; NULL: use g
    VertexId       *out_queue;              // optional: also enqueue found vertices
    unsigned long long out_cap;
    LevelCounters  *slots;
    long long       queue_index;
    long long       iteration;
};

template <typename Policy, typename Problem, typename Functor>
__global__ void __launch_bounds__(Policy::THREADS, Policy::MIN_CTAS)
Kernel(const __grid_constant__ typename Problem::DataSlice ds,
       const __grid_constant__ PullArgs<typename Policy::VertexId, typename Policy::SizeT> args)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    constexpr int THREADS = Policy::THREADS, WARPS = Policy::WARPS;
    constexpr int SERIAL_LIMIT = 8;
    constexpr bool NEED_SRC = Problem::MARK_PREDECESSORS;

    __shared__ unsigned long long s_red[THREADS / 32 + 1];

    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (blockIdx.x == 0 && threadIdx.x == 0) ClearFutureSlot(args.slots, args.queue_index);

    const int lane = util::LaneId();
    const long long local_nodes = args.g.local_nodes;
    const long long num_words = (local_nodes + 31) >> 5;
    // first word of this device's segment inside the global bitmaps
    const long long seg = ((long long)args.g.part.rank * local_nodes) >> 5;
    const long long warps_total = (long long)gridDim.x * WARPS;

    unsigned long long my_edges = 0, my_found = 0, my_degree = 0, my_cands = 0;

    for (long long w = (long long)blockIdx.x * WARPS + util::WarpId(); w < num_words; w += warps_total) {
        const unsigned vis = args.visited[seg + w];
        unsigned todo = ~vis;
        if (args.candidates) todo &= util::LoadStream(args.candidates + w);
        if (w == num_words - 1 && (local_nodes & 31)) todo &= (1u << (local_nodes & 31)) - 1;
        if (todo == 0) {                                 // uniform across the warp
            if (lane == 0) args.frontier_out[seg + w] = 0;
            continue;
        }
        const bool active = (todo >> lane) & 1u;
        const long long row = (w << 5) + lane;
        long long k = 0, end = 0, begin = 0;
        if (active) {
            begin = (long long)args.g.row_offsets[row];
            end   = (long long)args.g.row_offsets[row + 1];
            k = begin;
        }
        if (active) ++my_cands;
        bool found = false;
        VertexId parent = -1;
        long long epos = 0;

        // ---- lane-serial walk over the head of the list, early exit -----------------------
        {
            const long long lim = (end - begin > SERIAL_LIMIT) ? begin + SERIAL_LIMIT : end;
            while (k < lim) {
                const VertexId u = util::LoadStream(args.g.col_indices + k);
                if (util::BitTest(args.frontier_in, ds.BitPos(u))) { found = true; parent = u; epos = k; ++k; break; }
                ++k;
            }
            my_edges += (unsigned long long)(k - begin);
        }
        // ---- warp-cooperative walk over what is left of long lists -----------------------------
        unsigned rest = __ballot_sync(util::kFullMask, active && !found && k < end);
        while (rest) {
            const int leader = __ffs(rest) - 1;
            rest &= rest - 1;
            const long long wb = __shfl_sync(util::kFullMask, k, leader);
            const long long we = __shfl_sync(util::kFullMask, end, leader);
            for (long long p = wb; p < we; p += 32) {
                const long long idx = p + lane;
                VertexId u = -1;
                bool hit = false;
                if (idx < we) {
                    u = util::LoadStream(args.g.col_indices + idx);
                    hit = util::BitTest(args.frontier_in, ds.BitPos(u));
                    ++my_edges;
                }
                const unsigned hm = __ballot_sync(util::kFullMask, hit);
                if (hm) {
                    const int first = __ffs(hm) - 1;
                    const VertexId pu = __shfl_sync(util::kFullMask, u, first);
                    if (lane == leader) { found = true; parent = pu; epos = p + first; }
                    break;
                }
            }
        }
        // ---- functor + publication ----------------------------------------------------------
        const VertexId child = (VertexId)args.g.part.Global(row);
        if (found) {
            const VertexId s_id = NEED_SRC ? parent : (VertexId)args.iteration;
            found = VisitEdge<Policy, Problem, Functor>(&ds, s_id, child, (SizeT)epos, (SizeT)row);
        }
        const unsigned fm = __ballot_sync(util::kFullMask, found);
        if (lane == 0) {
            args.frontier_out[seg + w] = fm;
            if (fm) args.visited[seg + w] = vis | fm;
            my_found += __popc(fm);
        }
        if (found) {
            if (args.fwd_row_offsets)
                my_degree += (unsigned long long)(args.fwd_row_offsets[row + 1] - args.fwd_row_offsets[row]);
            else
                my_degree += (unsigned long long)(end - begin);
        }
        if (args.out_queue && fm) {
            unsigned long long b = 0;
            if (lane == 0) b = atomicAdd(&slot_out->out_len, (unsigned long long)__popc(fm));
            b = __shfl_sync(util::kFullMask, b, 0);
            if (found) {
                const unsigned long long at = b + __popc(fm & ((1u << lane) - 1));
                if (at < args.out_cap) args.out_queue[at] = child;
                else slot_out->overflow = 1;
            }
        }
    }
    BlockAccumulate<THREADS>(my_edges, s_red, &slot_out->edges);
    BlockAccumulate<THREADS>(my_degree, s_red, &slot_out->out_degree);
    BlockAccumulate<THREADS>(my_cands, s_red, &slot_out->aux0);
    // found vertices are counted in out_len unless the queue path already did
    if (!args.out_queue) BlockAccumulate<THREADS>(my_found, s_red, &slot_out->out_len);
}

}  // namespace pull
}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
