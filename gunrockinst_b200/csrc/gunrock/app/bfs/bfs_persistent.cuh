// bfs_persistent.cuh -- the whole (direction-optimising) BFS level loop as ONE cooperative kernel.
//
// The reference's enactors (gunrock/app/bfs/bfs_enactor.cuh:208-553, app/dobfs/
// dobfs_enactor.cuh:207-747) drive every level from the host: 4-6 launches, two blocking
// 4-byte memcpys and a mapped flag per level, and the direction decision on the CPU.  On a
// B200 an R-MAT scale-22 level is 5-50 us of device work, so those round trips (~20 us each)
// are most of the search.  Here the loop itself runs on the device:
//
//   148 x k co-resident CTAs (cooperative launch) execute, per level, the SAME operator
//   bodies the stand-alone kernels wrap -- advance::twc::ExpandItems, advance::lb::ScanTiles +
//   ExpandTiles, advance::pull::Sweep -- separated by grid-wide barriers.  After the barrier
//   every thread reads the level's 64-byte counter slot and takes the same decision
//   (terminate / TWC or LB / Beamer's push<->pull switch), so there is no host round trip,
//   no relaunch and no second copy of the policy.  The host launches once and fetches one
//   summary (+ optional per-level trace) at the end.
//
// Memory-model notes.  Data written in one phase and read in the next by other SMs crosses a
// cooperative-groups grid.sync(), which carries a gpu-scope fence (and, on sm_100, invalidates
// L1).  Frontier queues and the visited words are read with ld.global.cg (L2), bitmaps that are
// probed at random are read through L1 (ld.global.ca) and are never written in the phase that
// probes them; the CSR arrays are immutable and use the non-coherent path.
#pragma once
#include <cooperative_groups.h>
#include "bfs_functor.cuh"
#include "../enactor_base.cuh"

namespace gunrock {
namespace app {
namespace bfs {

// what one level did (device-written; the host turns it into LevelRecord / EnactorStats)
struct DeviceLevelRecord {
    int level, kind;                                // kind: 0 push-TWC, 1 pull, 2 push-LB, 3 push-LB on a CTA-private scan
    unsigned long long frontier, edges, discovered, candidates;
    unsigned long long ns;                          // globaltimer span of the level, barriers included
    unsigned long long ns_switch;                   // ... of which queue <-> bitmap conversion (+ its barrier)
};

struct PersistentSummary {
    unsigned long long levels, push_levels, pull_levels;
    unsigned long long total_queued, edges_inspected, edges_visited, algo_bytes;
    unsigned long long lb_calls;                    // LB scans of this search
    unsigned long long epoch_end;                   // look-back epoch of the LB scratch after this search
    unsigned long long overflow;
    unsigned long long grid_syncs;
    unsigned long long ns_kernel;                   // globaltimer span of the whole kernel as seen by its first thread
    unsigned long long ns_reset;                    // ... of which the in-kernel Problem::Reset
};

template <typename VertexId, typename SizeT>
struct PersistentParams {
    graph::CsrView<VertexId, SizeT> g;              // forward graph (push)
    graph::CsrView<VertexId, SizeT> ginv;           // inverse graph (pull); == g when symmetric
    const SizeT *fwd_row_offsets;                   // out-degrees for pull discoveries; NULL: ginv's
    VertexId    *queues[2];
    unsigned long long queue_cap[2];
    unsigned    *frontier_map[2];
    unsigned    *visited;
    const unsigned *candidates;
    const VertexId *head;
    unsigned long long *tickets;                    // 2 sets of pull-sweep chunk counters (zeroed by the host)
    oprtr::advance::LbScratchView<SizeT> scratch;   // .epoch = first epoch this run may use
    oprtr::LevelCounters *slots;                    // zeroed by the host before the launch
    DeviceLevelRecord    *records;                  // optional trace
    int                   max_records;
    PersistentSummary    *summary;
    VertexId src;                                   // the source, or ...
    const VertexId *sources;                        // ... num_sources of them, searched one after the other by ONE
    int      num_sources;                           //     launch (needs do_reset); summary[i] describes search i
    const VertexId *src_map;                        // optional: sources[] are ids of the caller's graph, the graph
                                                    //     here is its degree-ordered twin (src_map[v] = id here)
    const VertexId *to_new, *to_old;                // twin: every search ends by gathering its labels (and predecessors)
    VertexId *out_labels, *out_preds;               //     back into the caller's ids (NULL: not a twin)
    int      do_reset;                              // 1: the kernel also does Problem::Reset's device work
    unsigned long long mask_words;                  // words of the visited bitmap (do_reset)
    int      can_pull;
    int      traversal_mode;
    double   alpha, beta;
    unsigned long long lb_min_edges;
};

__device__ __forceinline__ unsigned long long GlobalTimerNs()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

template <typename PushPolicy, typename PullPolicy, bool NEED_SRC>
union PersistentStorage {
    // (the CTA-private scan of a short frontier is only alive together with the expansion: keeping it
    //  inside the union keeps the kernel's shared-memory carve-out, and with it the L1 left for the
    //  pull sweep's frontier-bit probes, close to the stand-alone kernels')
    struct {
        oprtr::advance::lb::ExpandStorage<PushPolicy, NEED_SRC> expand;
        oprtr::advance::lb::SmallScanStorage<PushPolicy>        small;
    } push;
    oprtr::advance::lb::ScanStorage<PushPolicy::THREADS>    scan;
    oprtr::advance::twc::Storage<PushPolicy, NEED_SRC>      twc;
    oprtr::advance::pull::SweepStorage<PullPolicy>          sweep;
    struct { unsigned warp[PushPolicy::THREADS / 32 + 1]; unsigned long long base; } b2q;
};

// frontier queue -> frontier bitmap (the bitmap is all-zero on entry)
template <typename VertexId, typename DataSlice>
__device__ __forceinline__ void QueueToBitmap(const VertexId *queue, unsigned long long len, unsigned *bitmap,
                                              const DataSlice &ds, unsigned long long tid, unsigned long long nthreads)
{
    for (unsigned long long i = tid; i < len; i += nthreads) {
        const VertexId v = util::LoadCG(queue + i);
        if (v >= 0) {
            const unsigned long long pos = ds.BitPos(v);
            atomicOr(bitmap + (pos >> 5), 1u << (pos & 31));
        }
    }
}

// this device's words of the frontier bitmap -> frontier queue; the words are cleared on the way
// (invariant of the persistent loop: the current map is all-zero while the search pushes)
template <int THREADS, typename VertexId, typename DataSlice, typename Shared>
__device__ __forceinline__ void BitmapToQueue(unsigned *bitmap, VertexId *queue, unsigned long long cap,
                                              unsigned long long *cursor, unsigned long long *overflow,
                                              const DataSlice &ds, Shared &sm, unsigned cta, unsigned num_ctas)
{
    const long long words = (ds.local_nodes + 31) >> 5;
    const long long seg = ((long long)ds.part.rank * ds.local_nodes) >> 5;
    const long long rounds = (words + (long long)num_ctas * THREADS - 1) / ((long long)num_ctas * THREADS);
    for (long long r = 0; r < rounds; ++r) {
        const long long w = (r * num_ctas + cta) * THREADS + threadIdx.x;
        unsigned bits = 0;
        if (w < words) {
            bits = util::LoadCG(bitmap + seg + w);
            if (bits) bitmap[seg + w] = 0;
        }
        unsigned total;
        const unsigned excl = util::BlockExclusiveSum<THREADS>((unsigned)__popc(bits), sm.warp, total);
        if (total == 0) continue;                               // uniform
        if (threadIdx.x == 0) sm.base = atomicAdd(cursor, (unsigned long long)total);
        __syncthreads();
        unsigned long long at = sm.base + excl;
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            if (at < cap) queue[at] = (VertexId)ds.part.Global((w << 5) + b);
            else *overflow = 1;
            ++at;
        }
        __syncthreads();
    }
}

#ifndef GRB_PERSISTENT_SPECIALISE
#define GRB_PERSISTENT_SPECIALISE true
#endif
#ifndef GRB_SWEEP_CALL
#define GRB_SWEEP_CALL __forceinline__
#endif
#ifndef GRB_PERSISTENT_CTAS
#define GRB_PERSISTENT_CTAS 4
#endif
constexpr int kPersistentCtasPerSm = GRB_PERSISTENT_CTAS;

// The pull sweep as a real call: inlined into the level loop it shares the 64-register budget with
// every value the loop keeps alive and spills (~0.85 KB per thread; the sweep of a scale-25 level
// ran 1.7x slower than the stand-alone pull::Kernel); as a callee it gets the whole budget and
// only the loop state that is live across the call is saved.
template <typename PullPolicy, typename Problem, typename PullFunctor>
__device__ GRB_SWEEP_CALL void SweepCall(
    const typename Problem::DataSlice &ds,
    const oprtr::advance::pull::PullArgs<typename Problem::VertexId, typename Problem::SizeT> &pa,
    oprtr::advance::pull::SweepStorage<PullPolicy> &sm, oprtr::LevelCounters *slot_out,
    long long first_chunk, long long chunk_stride)
{
    oprtr::advance::pull::Sweep<PullPolicy, Problem, PullFunctor, GRB_PERSISTENT_SPECIALISE>(ds, pa, sm, slot_out, first_chunk, chunk_stride);
}

template <typename PushPolicy, typename PullPolicy, typename Problem, typename PushFunctor, typename PullFunctor>
__global__ void __launch_bounds__(PushPolicy::THREADS, kPersistentCtasPerSm)
PersistentKernel(const __grid_constant__ typename Problem::DataSlice ds0,
                 const __grid_constant__ PersistentParams<typename Problem::VertexId, typename Problem::SizeT> P,
                 unsigned long long *cursor, unsigned long long desc_words)
{
    namespace cg  = cooperative_groups;
    namespace adv = oprtr::advance;
    typedef typename Problem::VertexId  VertexId;
    typedef typename Problem::SizeT     SizeT;
    typedef typename Problem::DataSlice DataSlice;
    constexpr int THREADS = PushPolicy::THREADS;
    static_assert(PushPolicy::THREADS == PullPolicy::THREADS, "one CTA shape for every phase");

    __shared__ PersistentStorage<PushPolicy, PullPolicy, Problem::MARK_PREDECESSORS> sm;
    __shared__ unsigned long long s_level[5];           // the level's counters: one L2 read per CTA, not per warp
    cg::grid_group grid = cg::this_grid();
    const unsigned cta = blockIdx.x, num_ctas = gridDim.x;
    const bool leader = cta == 0 && threadIdx.x == 0;
    const unsigned long long gtid = (unsigned long long)cta * THREADS + threadIdx.x;
    const unsigned long long gthreads = (unsigned long long)num_ctas * THREADS;

    const unsigned long long n = (unsigned long long)P.g.nodes;
    const unsigned osz = (unsigned)sizeof(SizeT);
    unsigned epoch = P.scratch.epoch;
    DataSlice ds = ds0;
    const int num_searches = P.sources ? P.num_sources : 1;

  for (int si = 0; si < num_searches; ++si) {
    const unsigned long long t_start = leader ? GlobalTimerNs() : 0ull;
    // ---- loop state: identical in every thread of the grid ------------------------------------
    const VertexId src = P.sources ? (P.src_map ? P.src_map[P.sources[si]] : P.sources[si]) : P.src;
    const bool src_is_mine = P.g.part.Mine(src);
    unsigned long long F = src_is_mine ? 1ull : 0ull;
    unsigned long long frontier_edges = 0;
    if (src_is_mine) {
        const long long row = P.g.part.Local(src);
        frontier_edges = (unsigned long long)(P.g.row_offsets[row + 1] - P.g.row_offsets[row]);
    }
    unsigned long long unexplored = (unsigned long long)P.g.edges - frontier_edges;
    unsigned long long prev_len = 0;
    bool pull = false, aborted = false;
    int  cur_map = 0, sel = 0;
    long long qi = 0;
    VertexId level = 0;
    PersistentSummary sum;
    sum.levels = sum.push_levels = sum.pull_levels = 0;
    sum.total_queued = sum.edges_inspected = sum.algo_bytes = 0;
    sum.edges_visited = frontier_edges;
    sum.lb_calls = sum.overflow = sum.grid_syncs = 0;

    // map 0 has to be all-zero while the search pushes; the barrier that ends level 0 covers this
    if (P.can_pull) {
        const unsigned long long words = (n + 31) / 32 + 1;
        for (unsigned long long w = gtid; w < words; w += gthreads) P.frontier_map[0][w] = 0;
    }
    // ---- BFSProblem::Reset on the device (reference: bfs_problem.cuh:272-357): labels <- -1,
    // preds <- -2, mask <- {src}, queue[0] = {src}, labels[src] = 0, preds[src] = -1; every element
    // is written once, by the thread that owns it, so no ordering between the fills is needed
    if (P.do_reset) {
        const long long ln = ds.local_nodes;
        const long long src_row = src_is_mine ? (long long)P.g.part.Local(src) : -1;
        const long long quads = ln >> 2;
        int4 *labels4 = reinterpret_cast<int4 *>(ds.d_labels);
        for (long long q = (long long)gtid; q < quads; q += (long long)gthreads) {
            int4 v = make_int4(-1, -1, -1, -1);
            if ((src_row >> 2) == q) {
                const int k = (int)(src_row & 3);
                if (k == 0) v.x = 0; else if (k == 1) v.y = 0; else if (k == 2) v.z = 0; else v.w = 0;
            }
            labels4[q] = v;
        }
        for (long long i = (quads << 2) + (long long)gtid; i < ln; i += (long long)gthreads)
            ds.d_labels[i] = (i == src_row) ? 0 : -1;
        if constexpr (Problem::MARK_PREDECESSORS) {
            for (long long i = (long long)gtid; i < ln; i += (long long)gthreads)
                ds.d_preds[i] = (i == src_row) ? -1 : -2;
        }
        const unsigned long long src_pos = ds.BitPos(src);
        for (unsigned long long w = gtid; w < P.mask_words; w += gthreads)
            P.visited[w] = (w == (src_pos >> 5)) ? (1u << (src_pos & 31)) : 0u;
        if (leader) {
            if (src_is_mine) P.queues[0][0] = src;
            unsigned long long *z = reinterpret_cast<unsigned long long *>(P.slots);
            for (int i = 0; i < (int)(4 * sizeof(oprtr::LevelCounters) / 8); ++i) z[i] = 0;
            P.summary[si].overflow = 0;
            *cursor = 0;
        }
        if (cta == 0 && P.tickets)          // both sets of pull-sweep chunk counters
            for (int x = threadIdx.x; x < 2 * oprtr::WorkProgress::kTicketWords; x += THREADS) P.tickets[x] = 0;
        grid.sync(); ++sum.grid_syncs;
    }

    sum.ns_reset = leader ? GlobalTimerNs() - t_start : 0ull;
    while (F > 0) {
        const unsigned long long t0 = leader ? GlobalTimerNs() : 0ull;
        ds.level = level;
        // ---- direction (Beamer's rule, as in BFSEnactor::EnactFused) ---------------------------
        if (P.can_pull) {
            if (!pull && (double)frontier_edges > (double)unexplored / P.alpha) {
                QueueToBitmap(P.queues[sel], F, P.frontier_map[cur_map], ds, gtid, gthreads);
                grid.sync(); ++sum.grid_syncs;
                pull = true;
            } else if (pull && F < prev_len && (double)F < (double)n / P.beta) {
                BitmapToQueue<THREADS>(P.frontier_map[cur_map], P.queues[sel], P.queue_cap[sel], cursor,
                                       &P.summary[si].overflow, ds, sm.b2q, cta, num_ctas);
                grid.sync(); ++sum.grid_syncs;
                if (leader) *cursor = 0;
                pull = false;
            }
        }
        const unsigned long long t1 = leader ? GlobalTimerNs() : 0ull;
        // ---- one level ------------------------------------------------------------------------------
        oprtr::LevelCounters *slot_out = P.slots + ((qi + 1) & 3);
        if (leader) oprtr::ClearFutureSlot(P.slots, qi);
        // chunk-ticket counters of the pull sweep: a level uses set (level & 1); every level clears the
        // other set, so whichever level pulls next finds its set zeroed
        if (cta == 0 && P.tickets && threadIdx.x < oprtr::WorkProgress::kTicketWords)
            P.tickets[((level + 1) & 1) * oprtr::WorkProgress::kTicketWords + threadIdx.x] = 0;
        int kind;
        if (!pull) {
            adv::Args<VertexId, SizeT> a;
            a.g = P.g;
            a.in_queue = P.queues[sel]; a.out_queue = P.queues[sel ^ 1];
            a.in_len = F; a.d_in_len = nullptr;
            a.out_cap = P.queue_cap[sel ^ 1];
            a.slots = P.slots; a.queue_index = qi; a.iteration = level;
            a.output = adv::OUT_COMPACT;
            const bool lb = P.traversal_mode == 1 || (P.traversal_mode != 2 && frontier_edges >= P.lb_min_edges);
            if (P.traversal_mode == 0 && F <= (unsigned long long)adv::lb::SmallScanStorage<PushPolicy>::SMALL_ITEMS) {
                // short frontier (first / last levels, a hub on its own): every CTA scans it privately
                // and expands its share of the edge tiles -- one pass, all SMs, no extra barrier
                kind = 3;
                adv::lb::ExpandSmall<PushPolicy, Problem, PushFunctor, GRB_PERSISTENT_SPECIALISE>(ds, a, sm.push.small, sm.push.expand, F, frontier_edges,
                                                                       slot_out, cta, num_ctas);
            } else if (!lb) {
                kind = 0;
                adv::twc::ExpandItems<PushPolicy, Problem, PushFunctor>(ds, a, sm.twc, F, slot_out, cta, num_ctas);
            } else {
                kind = 2;
                if (epoch + 1 >= adv::lb::kEpochLimit) {        // descriptor tags wrapped: start over
                    for (unsigned long long w = gtid; w < desc_words; w += gthreads) P.scratch.tile_desc[w] = 0;
                    grid.sync(); ++sum.grid_syncs;
                    epoch = 0;
                }
                ++epoch; ++sum.lb_calls;
                adv::LbScratchView<SizeT> sv = P.scratch;
                sv.epoch = epoch;
                adv::lb::ScanTiles<PushPolicy, THREADS, 4>(a, sv, sm.scan, F, slot_out, true);
                grid.sync(); ++sum.grid_syncs;
                if (leader) *sv.ticket = 0;
                adv::lb::ExpandTiles<PushPolicy, Problem, PushFunctor, GRB_PERSISTENT_SPECIALISE>(ds, a, sv, sm.push.expand, F, slot_out, cta, num_ctas);
            }
        } else {
            kind = 1;
            adv::pull::PullArgs<VertexId, SizeT> pa;
            pa.g = P.ginv;
            pa.frontier_in  = P.frontier_map[cur_map];
            pa.frontier_out = P.frontier_map[cur_map ^ 1];
            pa.visited = P.visited;
            pa.candidates = P.candidates;
            pa.head = P.head;
            pa.tickets = P.tickets + (level & 1) * oprtr::WorkProgress::kTicketWords;
            pa.fwd_row_offsets = P.fwd_row_offsets;
            pa.out_queue = nullptr; pa.out_cap = 0;
            pa.slots = P.slots; pa.queue_index = qi; pa.iteration = level;
            SweepCall<PullPolicy, Problem, PullFunctor>(
                ds, pa, sm.sweep, slot_out, (long long)cta * PullPolicy::WARPS + util::WarpId(),
                (long long)num_ctas * PullPolicy::WARPS);
        }
        grid.sync(); ++sum.grid_syncs;
        // ---- every thread reads the level's counters and updates the same state ----------------
        if (threadIdx.x == 0) {
            s_level[0] = util::LoadCG(&slot_out->out_len);
            s_level[1] = util::LoadCG(&slot_out->edges);
            s_level[2] = util::LoadCG(&slot_out->out_degree);
            s_level[3] = util::LoadCG(&slot_out->aux0);
            s_level[4] = util::LoadCG(&slot_out->overflow);
        }
        __syncthreads();
        const unsigned long long D = s_level[0], E = s_level[1], deg = s_level[2], cands = s_level[3], ovf = s_level[4];
        __syncthreads();                                        // s_level is rewritten after the next level
        sum.total_queued += F;
        sum.edges_inspected += E;
        if (!pull) { ++sum.push_levels; sum.algo_bytes += F * (4 + 2 * osz) + 5 * E + 8 * D; }
        else       { ++sum.pull_levels; sum.algo_bytes += cands * (2 * osz) + 5 * E + 4 * D + 3 * (n / 8); }
        if (leader && (int)sum.levels < P.max_records) {
            DeviceLevelRecord r;
            r.level = level; r.kind = kind; r.frontier = F; r.edges = E; r.discovered = D; r.candidates = cands;
            r.ns = GlobalTimerNs() - t0;
            r.ns_switch = t1 - t0;
            P.records[sum.levels] = r;
        }
        ++sum.levels;
        if (ovf) { sum.overflow = 1; aborted = true; break; }
        prev_len = F;
        F = D;
        frontier_edges = deg;
        unexplored = unexplored > frontier_edges ? unexplored - frontier_edges : 0;
        sum.edges_visited += frontier_edges;
        ++qi;
        if (pull) cur_map ^= 1; else sel ^= 1;
        ++level;
    }
    // degree-ordered twin: the search's result in the caller's vertex ids (every label store of the last
    // level is behind a grid-wide barrier; the next search's reset is behind the one that ends this search)
    if (P.out_labels && !aborted) {
        for (unsigned long long v = gtid; v < n; v += gthreads) {
            const VertexId t = P.to_new[v];
            P.out_labels[v] = util::LoadCG(ds.d_labels + t);
            if constexpr (Problem::MARK_PREDECESSORS) {
                const VertexId p = util::LoadCG(ds.d_preds + t);
                P.out_preds[v] = p >= 0 ? P.to_old[p] : p;
            }
        }
    }
    if (leader) {
        if (P.summary[si].overflow) sum.overflow = 1;  // raised by the bitmap -> queue conversion
        sum.ns_kernel = GlobalTimerNs() - t_start;
        sum.epoch_end = epoch;
        P.summary[si] = sum;
    }
    if (aborted) break;                                 // uniform: every thread saw the same counter
    // the next search resets the counters every CTA has just read: everybody must be past that read
    if (si + 1 < num_searches) grid.sync();
  }
}

}  // namespace bfs
}  // namespace app
}  // namespace gunrock
