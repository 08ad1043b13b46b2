// bfs_enactor.cuh -- the host loop of (direction-optimising) BFS.
//
// Counterpart of the reference's BFSEnactor::EnactBFS (gunrock/app/bfs/
// bfs_enactor.cuh:208-553) and DOBFSEnactor::EnactDOBFS (app/dobfs/
// dobfs_enactor.cuh:207-747).  Per level the reference issues advance (4-5
// launches in LB mode) + filter and crosses the PCIe bus three times (two
// blocking 4-byte memcpys and a mapped flag), and this
// fork adds an RPC + relaunch loop; its DOBFS needs four extra filter passes
// to move between queue and bool-map form.
//
// Here a level is
//     push:  [ScanKernel +] ExpandKernel  or  twc::Kernel     (fused advance + filter)
//     pull:  pull::Kernel                                     (bitmap sweep)
// followed by ONE 64-byte counter fetch that tells the host the next frontier
// length, its degree sum (sizes the next grid and drives the direction switch)
// and the work done.  Direction switches cost one bitmap<->queue kernel.
//
// Direction heuristic: Beamer's edge-based rule -- go bottom-up when the
// frontier's out-edges exceed (unexplored edges)/alpha, return top-down when
// the frontier is shrinking and smaller than n/beta.  (The reference's rule
// compares vertex counts, dobfs_enactor.cuh:397,569, with alpha=12, beta=6 from
// test_dobfs.cu:533-534; depths do not depend on the rule.)
#pragma once
#ifndef GRB_FILTER_IPT
#define GRB_FILTER_IPT 4      // filter: items per thread per tile
#endif
#ifndef GRB_FILTER_CTAS
#define GRB_FILTER_CTAS 6     // filter: resident CTAs per SM
#endif
#ifndef GRB_LB_CTAS
#define GRB_LB_CTAS 8        // resident CTAs per SM of the push-direction advance kernels
#endif
#ifndef GRB_PULL_CTAS
#define GRB_PULL_CTAS 6      // resident CTAs per SM the stand-alone pull sweep is bounded for
#endif
#include <vector>
#include <cstdlib>
#include <cstdio>
#include "bfs_functor.cuh"
#include "../enactor_base.cuh"
#include "../../graph/build.cuh"
#include "bfs_persistent.cuh"

namespace gunrock {
namespace app {
namespace bfs {

// frontier queue -> frontier bitmap (bitmap zeroed by the caller)
template <typename VertexId, typename DataSlice>
__global__ void QueueToBitmapKernel(const VertexId *queue, unsigned long long len, unsigned *bitmap, DataSlice ds)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += stride) {
        const VertexId v = queue[i];
        if (v >= 0) {
            const unsigned long long pos = ds.BitPos(v);
            atomicOr(bitmap + (pos >> 5), 1u << (pos & 31));
        }
    }
}

// this device's words of the frontier bitmap -> frontier queue (CTA-aggregated enqueue)
template <typename VertexId, typename DataSlice, int THREADS>
__global__ void BitmapToQueueKernel(const unsigned *bitmap, VertexId *queue, unsigned long long cap,
                                    oprtr::LevelCounters *slot, DataSlice ds)
{
    __shared__ unsigned s_warp[THREADS / 32 + 1];
    __shared__ unsigned long long s_base;
    const long long words = (ds.local_nodes + 31) >> 5;
    const long long seg = ((long long)ds.part.rank * ds.local_nodes) >> 5;
    const long long rounds = (words + (long long)gridDim.x * THREADS - 1) / ((long long)gridDim.x * THREADS);
    for (long long r = 0; r < rounds; ++r) {
        const long long w = (r * gridDim.x + blockIdx.x) * THREADS + threadIdx.x;
        unsigned bits = (w < words) ? bitmap[seg + w] : 0u;
        unsigned total;
        unsigned excl = util::BlockExclusiveSum<THREADS>((unsigned)__popc(bits), s_warp, total);
        if (total == 0) continue;
        if (threadIdx.x == 0) s_base = atomicAdd(&slot->out_len, (unsigned long long)total);
        __syncthreads();
        unsigned long long at = s_base + excl;
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            if (at < cap) queue[at] = (VertexId)ds.part.Global((w << 5) + b);
            ++at;
        }
        __syncthreads();
    }
}

// sum of out-degrees of labelled vertices: the numerator of the reference's MiEdges/s figure,
// which it computes on the host after Extract (gunrock/app/bfs/bfs_app.cu:82-91)
template <typename VertexId, typename SizeT>
__global__ void EdgesVisitedKernel(const VertexId *labels, const SizeT *row_offsets, long long local_nodes,
                                   unsigned long long *out)
{
    __shared__ unsigned long long s_red[256 / 32 + 1];
    unsigned long long mine = 0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < local_nodes; v += stride)
        if (labels[v] > -1) mine += (unsigned long long)(row_offsets[v + 1] - row_offsets[v]);
    oprtr::advance::BlockAccumulate<256>(mine, s_red, out);
}

struct BFSOptions {
    int    traversal_mode      = 0;      // 0: choose per level, 1: always LB, 2: always TWC (reference: --traversal-mode)
    bool   direction_optimized = false;  // allow the pull direction (needs an inverse graph)
    double alpha               = 8.0;    // push -> pull when frontier_edges > unexplored_edges / alpha
    double beta                = 72.0;   // pull -> push when frontier < n / beta and shrinking (swept on B200: tools/sweep_alpha.py)
    unsigned long long lb_min_edges = 1u << 12;   // mode 0: LB from this many frontier edges upward
    bool   record_levels       = false;  // keep a per-level trace (adds two events per level)
    bool   host_loop           = false;  // drive every level from the host (the reference's structure) instead
                                         // of the single cooperative kernel of bfs_persistent.cuh
    // host loop: push levels with at least this many frontier edges (idempotent mode: frontier items) take the
    // streaming advance (oprtr/advance/lb_stream.cuh); 0 = never
    unsigned long long stream_min_edges = 1ull << 20;
};
// (the runner leaves stream_min_edges at its default only for degree-ordered twins: with the caller's vertex order the hot
//  range of the visited map answers few probes and lb::ExpandKernel / filter::Kernel are as fast or faster -- scale 25,
//  top-down: 8.5 against 7.2 ms; on the twin 4.2 against 6.2 ms)

struct LevelRecord {
    int level; int pull;                 // pull: 0 push-TWC, 1 pull, 2 push-LB, 5 push-LB streaming (lb_stream.cuh)
    unsigned long long frontier, edges, discovered, candidates;
    float ms;
    float switch_ms;                     // part of ms spent converting queue <-> bitmap (persistent loop)
};

template <typename Problem>
class BFSEnactor : public EnactorBase {
  public:
    typedef typename Problem::VertexId VertexId;
    typedef typename Problem::SizeT    SizeT;
    typedef typename Problem::Value    Value;
    typedef typename Problem::DataSlice DataSlice;
    typedef BFSFunctor<VertexId, SizeT, Value, Problem>        BfsFunctor;
    typedef ReverseBFSFunctor<VertexId, SizeT, Value, Problem> PullFunctor;

    // sm_100a policies: 256-thread CTAs, 4 neighbour entries per thread per tile,
    // grids of 148 x MIN_CTAS persistent CTAs; operator-level bitmask pre-cull on,
    // degree tracking on for the fused (non-idempotent) mode
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, GRB_LB_CTAS, true, true>   AdvancePolicy;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, true, false>  IdemAdvancePolicy;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, GRB_PULL_CTAS, false, false> PullPolicy;
    // streaming advance (lb_stream.cuh): one 1024-thread CTA per SM, 8 producer + 24 consumer warps, 6144-entry tiles
    typedef oprtr::advance::KernelPolicy<Problem, 1024, 6, 1, true, true>   StreamPolicy;
    typedef oprtr::advance::KernelPolicy<Problem, 1024, 6, 1, true, false>  IdemStreamPolicy;   // the filter's streaming instance
    // reference: END_BITMASK_CULL = 5 (bfs_enactor.cuh:594); the word-atomic mask stays exact, so keep it on
    typedef oprtr::filter::KernelPolicy<Problem, 256, GRB_FILTER_IPT, GRB_FILTER_CTAS, -1> FilterPolicy;

    oprtr::advance::LbScratch<SizeT> lb_scratch;
    std::vector<LevelRecord> levels;
    unsigned long long edges_visited = 0;       // sum of out-degrees of visited vertices (reference GTEPS numerator)
    cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr;
    bool set_up = false;
    // persistent (single cooperative kernel) loop
    static const int kMaxRecords = 1024;
    util::DeviceArray<DeviceLevelRecord>  d_records;
    util::DeviceArray<PersistentSummary>  d_summary;     // [0] summary, [1] reused as the 8-byte queue cursor
    util::PinnedArray<DeviceLevelRecord>  h_records;
    util::PinnedArray<PersistentSummary>  h_summary;
    util::DeviceArray<VertexId>           d_sources;     // run-many: the batch of sources
    util::PinnedArray<VertexId>           h_sources;
    int many_count = 0;
    bool persistent_ok = false;
    int  kernel_us = 0, reset_us = 0;     // device-side span of the last persistent search / of its reset phase
    int  persistent_grid = 0;

    ~BFSEnactor()
    {
        if (ev_a) cudaEventDestroy(ev_a);
        if (ev_b) cudaEventDestroy(ev_b);
        if (ev_c) cudaEventDestroy(ev_c);
    }

    // one-time allocations (outside any timed region; the reference cudaMallocs its
    // LB scratch inside every Enact, bfs_enactor.cuh:254-258)
    cudaError_t Setup(Problem *problem, cudaStream_t s)
    {
        GR_TRY(SetupBase(problem->graph->device, s));
        const size_t items = (size_t)problem->graph->local_nodes + 1;
        size_t max_items = Problem::ENABLE_IDEMPOTENCE ? (size_t)problem->frontier_elements[0] : items;
        if (max_items < items) max_items = items;
        GR_TRY(lb_scratch.Allocate(max_items, (size_t)problem->graph->edges + 1, AdvancePolicy::SCAN_TILE,
                                   AdvancePolicy::TILE_EDGES, s));
        if (!ev_a) { GR_TRY(cudaEventCreate(&ev_a)); GR_TRY(cudaEventCreate(&ev_b)); GR_TRY(cudaEventCreate(&ev_c)); }
        if constexpr (!Problem::ENABLE_IDEMPOTENCE) GR_TRY(SetupPersistent(problem));
        set_up = true;
        return cudaSuccess;
    }

    // reference signature: Enact<BFSProblem>(context, problem, src, max_grid_size, traversal_mode)
    // true when Enact will run the whole search, Problem::Reset's device work included, as one
    // kernel: the caller may then use Problem::Reset(..., on_device = true)
    bool FusesReset(VertexId src, const Problem *problem, const BFSOptions &opt) const
    {
        return set_up && !Problem::ENABLE_IDEMPOTENCE && persistent_ok && !opt.host_loop &&
               src >= 0 && (long long)src < problem->graph->nodes;
    }

    // sources / count / summaries: a batch of searches in one launch (needs reset_on_device; `src` is
    // then the last source of the batch, whose labels are the ones left on the device)
    cudaError_t Enact(Problem *problem, VertexId src, const BFSOptions &opt, bool reset_on_device = false,
                      const VertexId *sources = nullptr, int count = 0, PersistentSummary *summaries = nullptr)
    {
        if (!set_up) return util::GRError(cudaErrorNotReady, "BFSEnactor::Setup not called", __FILE__, __LINE__);
        if (reset_on_device && !FusesReset(src, problem, opt))
            return util::GRError(cudaErrorInvalidValue, "BFSEnactor: device-side reset needs the persistent kernel",
                                 __FILE__, __LINE__);
        if (reset_on_device) { enactor_stats = EnactorStats(); frontier_attribute = FrontierAttribute(); }
        else GR_TRY(ResetBase());
        levels.clear();
        edges_visited = 0;
        if (src < 0 || (long long)src >= problem->graph->nodes) return cudaSuccess;
        if constexpr (Problem::ENABLE_IDEMPOTENCE) {
            GR_TRY(EnactIdempotent(problem, opt));
            return GatherToCallerIds(problem);
        } else {
            if (sources && count > 0) {
                if (!reset_on_device)
                    return util::GRError(cudaErrorInvalidValue, "BFSEnactor: a batch needs the device-side reset",
                                         __FILE__, __LINE__);
                return EnactPersistentMany(problem, sources, count, opt, summaries);
            }
            if (opt.host_loop || !persistent_ok) {
                GR_TRY(EnactFused(problem, opt));
                return GatherToCallerIds(problem);
            }
            return EnactPersistent(problem, opt, reset_on_device);
        }
    }

    // degree-ordered twin, host-driven loops: the result in the caller's vertex ids (the persistent kernel
    // ends every search with the same gather)
    cudaError_t GatherToCallerIds(Problem *problem)
    {
        if (!problem->src_map) return cudaSuccess;
        const long long n = problem->graph->nodes;
        if (n == 0) return cudaSuccess;
        graph::GatherThroughKernel<false><<<util::FillGrid((unsigned long long)n), 256, 0, stream>>>(
            problem->labels.ptr, problem->src_map, problem->dst_map, n, problem->out_labels.ptr);
        if (Problem::MARK_PREDECESSORS)
            graph::GatherThroughKernel<true><<<util::FillGrid((unsigned long long)n), 256, 0, stream>>>(
                problem->preds.ptr, problem->src_map, problem->dst_map, n, problem->out_preds.ptr);
        ++enactor_stats.launches;
        return util::GRError(cudaGetLastError(), "GatherThroughKernel failed", __FILE__, __LINE__);
    }

    // recount edges_visited from the labels (idempotent mode cannot track it: duplicates)
    cudaError_t RecountEdgesVisited(Problem *problem)
    {
        oprtr::LevelCounters *slot = work_progress.Slot(0);
        GR_TRY(cudaMemsetAsync(slot, 0, sizeof(oprtr::LevelCounters), stream));
        const long long ln = problem->graph->local_nodes;
        EdgesVisitedKernel<VertexId, SizeT><<<util::FillGrid((unsigned long long)ln), 256, 0, stream>>>(
            problem->labels.ptr, problem->graph->row_offsets.ptr, ln, &slot->aux0);
        GR_TRY(cudaGetLastError());
        oprtr::LevelCounters c;
        GR_TRY(work_progress.Fetch(0, c, stream));
        edges_visited = c.aux0;
        return cudaSuccess;
    }

  private:
    typedef oprtr::advance::Args<VertexId, SizeT> AdvArgs;

    AdvArgs PushArgs(Problem *problem, VertexId level)
    {
        AdvArgs a;
        a.g = problem->graph->View();
        a.in_queue  = problem->frontier_queues[frontier_attribute.selector].ptr;
        a.out_queue = problem->frontier_queues[frontier_attribute.selector ^ 1].ptr;
        a.in_len = frontier_attribute.queue_length; a.d_in_len = nullptr;
        a.out_cap = problem->frontier_elements[frontier_attribute.selector ^ 1];
        a.slots = work_progress.d_slots.ptr;
        a.queue_index = frontier_attribute.queue_index;
        a.iteration = level;
        a.output = oprtr::advance::OUT_COMPACT;
        return a;
    }

    // words of the visited map the streaming advance mirrors in shared memory
    template <typename P>
    static unsigned HotWords(const Problem *problem)
    {
        const unsigned long long words = (unsigned long long)((problem->graph->nodes + 31) / 32);
        const unsigned cap = oprtr::advance::lb::StreamMaxHotWords<P>();
        return (unsigned)(words < cap ? words : cap);
    }

    oprtr::advance::MODE PickMode(const BFSOptions &opt, unsigned long long frontier_edges)
    {
        if (opt.traversal_mode == 1) return oprtr::advance::LB;
        if (opt.traversal_mode == 2) return oprtr::advance::TWC_FORWARD;
        return frontier_edges >= opt.lb_min_edges ? oprtr::advance::LB : oprtr::advance::TWC_FORWARD;
    }

    // ---- persistent mode: the level loop runs on the device (bfs_persistent.cuh) ------------------
    typedef PersistentParams<VertexId, SizeT> PParams;
    typedef void (*PersistentFn)(const DataSlice, const PParams, unsigned long long *, unsigned long long);
    static PersistentFn PersistentEntry()
    {
        return &PersistentKernel<AdvancePolicy, PullPolicy, Problem, BfsFunctor, PullFunctor>;
    }

    cudaError_t SetupPersistent(Problem *problem)
    {
        int coop = 0;
        GR_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev.device));
        int per_sm = 0;
        GR_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, PersistentEntry(), AdvancePolicy::THREADS, 0));
        if (per_sm > kPersistentCtasPerSm) per_sm = kPersistentCtasPerSm;
        persistent_ok = coop != 0 && per_sm > 0 && !getenv("GRB_NO_PERSISTENT");
        persistent_grid = per_sm * dev.sm_count;
        GR_TRY(d_records.Allocate(kMaxRecords));
        GR_TRY(d_summary.Allocate(2));
        GR_TRY(h_records.Allocate(kMaxRecords));
        GR_TRY(h_summary.Allocate(1));
        return cudaSuccess;
    }

    // `count` searches in one launch (sources on the host; needs the device-side reset); summaries[i]
    // receives the i-th search's summary
    cudaError_t EnactPersistentMany(Problem *problem, const VertexId *sources, int count, const BFSOptions &opt,
                                    PersistentSummary *summaries)
    {
        if (count <= 0) return cudaSuccess;
        GR_TRY(d_sources.Allocate((size_t)count));
        GR_TRY(h_sources.Allocate((size_t)count));
        GR_TRY(d_summary.Allocate((size_t)count + 1));
        GR_TRY(h_summary.Allocate((size_t)count + 1));
        for (int i = 0; i < count; ++i) h_sources.ptr[i] = sources[i];
        GR_TRY(cudaMemcpyAsync(d_sources.ptr, h_sources.ptr, sizeof(VertexId) * (size_t)count, cudaMemcpyHostToDevice, stream));
        problem->src = sources[count - 1];
        many_count = count;
        cudaError_t e = EnactPersistent(problem, opt, true);
        many_count = 0;
        for (int i = 0; i < count; ++i) summaries[i] = h_summary.ptr[i];
        return e;
    }

    cudaError_t EnactPersistent(Problem *problem, const BFSOptions &opt, bool reset_on_device)
    {
        typename Problem::Base::Graph &g = *problem->graph;
        const bool can_pull = opt.direction_optimized && g.HasInverse();
        if (can_pull) GR_TRY(problem->PreparePull());

        PParams P;
        P.g = g.View();
        P.ginv = g.InverseView();
        P.fwd_row_offsets = g.symmetric ? nullptr : g.row_offsets.ptr;
        for (int i = 0; i < 2; ++i) {
            P.queues[i] = problem->frontier_queues[i].ptr;
            P.queue_cap[i] = problem->frontier_elements[i];
            P.frontier_map[i] = can_pull ? problem->frontier_map[i].ptr : nullptr;
        }
        P.visited = problem->visited_mask.ptr;
        P.candidates = can_pull ? problem->candidates.ptr : nullptr;
        P.head = can_pull ? problem->head.ptr : nullptr;
        P.tickets = work_progress.d_tickets.ptr;
        GR_TRY(work_progress.d_tickets.Zero(stream));
        if (lb_scratch.epoch + 64 >= oprtr::advance::lb::kEpochLimit) {   // keep wraps out of the common case
            GR_TRY(lb_scratch.tile_desc.Zero(stream));
            lb_scratch.epoch = 0;
        }
        P.scratch.scanned = lb_scratch.scanned.ptr;     P.scratch.row_start = lb_scratch.row_start.ptr;
        P.scratch.tile_desc = lb_scratch.tile_desc.ptr; P.scratch.part_start = lb_scratch.part_start.ptr;
        P.scratch.ticket = lb_scratch.ticket.ptr;       P.scratch.part_cap = lb_scratch.part_start.cap;
        P.scratch.epoch = lb_scratch.epoch;
        P.slots = work_progress.d_slots.ptr;            // zeroed by ResetBase()
        P.records = d_records.ptr;
        P.max_records = opt.record_levels ? kMaxRecords : 0;
        P.summary = d_summary.ptr;
        P.src = problem->src;
        P.sources = many_count ? d_sources.ptr : nullptr;
        P.num_sources = many_count;
        P.src_map = many_count ? problem->src_map : nullptr;
        P.to_new = problem->src_map; P.to_old = problem->dst_map;
        P.out_labels = problem->src_map ? problem->out_labels.ptr : nullptr;
        P.out_preds = (problem->src_map && Problem::MARK_PREDECESSORS) ? problem->out_preds.ptr : nullptr;
        P.do_reset = reset_on_device ? 1 : 0;
        P.mask_words = (unsigned long long)problem->visited_mask.size;
        P.can_pull = can_pull ? 1 : 0;
        P.traversal_mode = opt.traversal_mode;
        P.alpha = opt.alpha; P.beta = opt.beta;
        P.lb_min_edges = opt.lb_min_edges;

        if (!reset_on_device || many_count) GR_TRY(d_summary.Zero(stream));
        DataSlice ds = problem->MakeDataSlice(0);
        const int searches = many_count ? many_count : 1;
        unsigned long long *cursor = reinterpret_cast<unsigned long long *>(d_summary.ptr + searches);
        unsigned long long desc_words = (unsigned long long)lb_scratch.tile_desc.size;
        void *kargs[] = {(void *)&ds, (void *)&P, (void *)&cursor, (void *)&desc_words};
        cudaError_t le = getenv("GRB_FAIL_COOP_LAUNCH") ? cudaErrorCooperativeLaunchTooLarge
                         : cudaLaunchCooperativeKernel((const void *)PersistentEntry(), dim3((unsigned)persistent_grid),
                                                       dim3(AdvancePolicy::THREADS), kargs, 0, stream);
        if (le != cudaSuccess) {
            // co-residency is not available after all (MPS / SM partitioning, another context, a changed
            // shared-memory carve-out): fall back to the host-driven loop for the rest of this context
            (void)cudaGetLastError();
            persistent_ok = false;
            fprintf(stderr, "libgunrock (B200 build): cooperative launch of the persistent BFS kernel failed (%s); "
                            "using the host-driven level loop\n", cudaGetErrorString(le));
            if (many_count) return cudaErrorCooperativeLaunchTooLarge;      // the caller re-runs the batch search by search
            if (reset_on_device) GR_TRY(problem->FillState(problem->src));
            GR_TRY(ResetBase());
            GR_TRY(EnactFused(problem, opt));
            return GatherToCallerIds(problem);
        }
        enactor_stats.launches = 1;
        GR_TRY(cudaMemcpyAsync(h_summary.ptr, d_summary.ptr, sizeof(PersistentSummary) * (size_t)searches,
                               cudaMemcpyDeviceToHost, stream));
        if (opt.record_levels)
            GR_TRY(cudaMemcpyAsync(h_records.ptr, d_records.ptr, sizeof(DeviceLevelRecord) * kMaxRecords,
                                   cudaMemcpyDeviceToHost, stream));
        GR_TRY(cudaStreamSynchronize(stream));
        // a search that overflowed its queues ends the batch: the later summaries stay zero.  Report the
        // overflow whichever search raised it, and take the look-back epoch from the last search that ran
        // (stale descriptor tags of a higher epoch must never look valid to a later scan).
        int last_ran = 0;
        bool overflowed = false;
        for (int i = 0; i < searches; ++i) {
            if (h_summary.ptr[i].ns_kernel != 0 || h_summary.ptr[i].levels != 0) last_ran = i;
            if (h_summary.ptr[i].overflow) overflowed = true;
        }
        const PersistentSummary &sum = h_summary.ptr[overflowed ? last_ran : searches - 1];   // stats of the (last) search
        lb_scratch.epoch = (unsigned)h_summary.ptr[last_ran].epoch_end;   // the kernel wraps (and re-zeroes) on its own
        if (overflowed) {
            GR_TRY(lb_scratch.tile_desc.Zero(stream));
            lb_scratch.epoch = 0;
        }
        enactor_stats.iteration = (long long)sum.levels;
        enactor_stats.total_queued = sum.total_queued;
        enactor_stats.edges_inspected = sum.edges_inspected;
        enactor_stats.algo_bytes = sum.algo_bytes;
        enactor_stats.push_levels = (int)sum.push_levels;
        enactor_stats.pull_levels = (int)sum.pull_levels;
        edges_visited = sum.edges_visited;
        kernel_us = (int)(sum.ns_kernel / 1000); reset_us = (int)(sum.ns_reset / 1000);
        if (opt.record_levels) {
            const int cnt = (int)(sum.levels < (unsigned long long)kMaxRecords ? sum.levels : kMaxRecords);
            for (int i = 0; i < cnt; ++i) {
                const DeviceLevelRecord &d = h_records.ptr[i];
                LevelRecord r;
                r.level = d.level; r.pull = d.kind; r.frontier = d.frontier; r.edges = d.edges;
                r.discovered = d.discovered; r.candidates = d.candidates; r.ms = (float)((double)d.ns * 1e-6);
                r.switch_ms = (float)((double)d.ns_switch * 1e-6);
                levels.push_back(r);
            }
        }
        if (overflowed)
            return util::GRError(cudaErrorInvalidConfiguration,
                                 "frontier queue overflow: increase queue_size", __FILE__, __LINE__);
        return cudaSuccess;
    }

    // ---- normal mode: fused advance + filter, optional direction optimisation -----------------
    cudaError_t EnactFused(Problem *problem, const BFSOptions &opt)
    {
        typename Problem::Base::Graph &g = *problem->graph;
        const bool can_pull = opt.direction_optimized && g.HasInverse();
        if (can_pull) GR_TRY(problem->PreparePull());

        const bool owner = g.part.Mine(problem->src);
        frontier_attribute.queue_length = owner ? 1 : 0;
        unsigned long long frontier_edges = problem->src_degree;
        unsigned long long unexplored = (unsigned long long)g.edges - frontier_edges;
        edges_visited = frontier_edges;
        unsigned long long prev_len = 0;
        bool pull = false;
        int  cur_map = 0;
        VertexId level = 0;
        const unsigned long long n = (unsigned long long)g.nodes;
        const unsigned osz = (unsigned)sizeof(SizeT);

        while (frontier_attribute.queue_length > 0) {
            const unsigned long long F = frontier_attribute.queue_length;
            // -- direction decision ---------------------------------------------------------------
            if (can_pull) {
                if (!pull && (double)frontier_edges > (double)unexplored / opt.alpha) {
                    GR_TRY(problem->frontier_map[cur_map].Zero(stream));
                    DataSlice ds = problem->MakeDataSlice(level);
                    QueueToBitmapKernel<VertexId, DataSlice><<<util::FillGrid(F), 256, 0, stream>>>(
                        problem->frontier_queues[frontier_attribute.selector].ptr, F, problem->frontier_map[cur_map].ptr, ds);
                    GR_TRY(cudaGetLastError());
                    ++enactor_stats.launches;
                    pull = true;
                } else if (pull && F < prev_len && (double)F < (double)n / opt.beta) {
                    DataSlice ds = problem->MakeDataSlice(level);
                    // borrow the slot of the upcoming kernel for the count, then clear it again
                    oprtr::LevelCounters *slot = work_progress.Slot(frontier_attribute.queue_index + 1);
                    BitmapToQueueKernel<VertexId, DataSlice, 256><<<util::FillGrid((n + 31) / 32), 256, 0, stream>>>(
                        problem->frontier_map[cur_map].ptr, problem->frontier_queues[frontier_attribute.selector].ptr,
                        problem->frontier_elements[frontier_attribute.selector], slot, ds);
                    GR_TRY(cudaGetLastError());
                    GR_TRY(cudaMemsetAsync(slot, 0, sizeof(oprtr::LevelCounters), stream));
                    ++enactor_stats.launches;
                    pull = false;
                }
            }
            // -- one level ------------------------------------------------------------------------------
            if (opt.record_levels) GR_TRY(cudaEventRecord(ev_a, stream));
            DataSlice ds = problem->MakeDataSlice(level);
            int kind;
            if (!pull) {
                oprtr::advance::MODE mode = PickMode(opt, frontier_edges);
                kind = (mode == oprtr::advance::LB) ? 2 : 0;
                enactor_stats.launches += (mode == oprtr::advance::LB) ? 2 : 1;
                bool streamed = false;
                if constexpr (!Problem::MARK_PREDECESSORS) {
                    AdvArgs a = PushArgs(problem, level);
                    if (mode == oprtr::advance::LB && opt.stream_min_edges && frontier_edges >= opt.stream_min_edges &&
                        (frontier_edges >= 8ull * F || opt.stream_min_edges <= 2) &&      // lists of ~1 entry: a warp per 32 items is latency
                        oprtr::advance::StreamEligible<StreamPolicy, Problem>(a)) {
                        kind = 5;
                        streamed = true;
                        ++enactor_stats.launches;          // scan + expand + harvest
                        GR_TRY((oprtr::advance::LaunchStreamKernel<StreamPolicy, Problem, BfsFunctor>(
                            dev, ds, a, &lb_scratch, frontier_edges, HotWords<StreamPolicy>(problem), stream)));
                    }
                }
                if (!streamed)
                    GR_TRY((oprtr::advance::LaunchKernel<AdvancePolicy, Problem, BfsFunctor>(
                        mode, dev, ds, PushArgs(problem, level), &lb_scratch, frontier_edges, stream)));
            } else {
                kind = 1;
                ++enactor_stats.launches;
                oprtr::advance::pull::PullArgs<VertexId, SizeT> pa;
                pa.g = g.InverseView();
                pa.frontier_in  = problem->frontier_map[cur_map].ptr;
                pa.frontier_out = problem->frontier_map[cur_map ^ 1].ptr;
                pa.visited      = problem->visited_mask.ptr;
                pa.candidates   = problem->candidates.ptr;
                pa.head         = problem->head.ptr;
                pa.tickets      = work_progress.Tickets(level);
                GR_TRY(cudaMemsetAsync(pa.tickets, 0, sizeof(unsigned long long) * oprtr::WorkProgress::kTicketWords, stream));
                pa.fwd_row_offsets = g.symmetric ? nullptr : g.row_offsets.ptr;
                pa.out_queue = nullptr; pa.out_cap = 0;
                pa.slots = work_progress.d_slots.ptr;
                pa.queue_index = frontier_attribute.queue_index;
                pa.iteration = level;
                GR_TRY((oprtr::advance::LaunchPullKernel<PullPolicy, Problem, PullFunctor>(dev, ds, pa, stream)));
            }
            if (opt.record_levels) GR_TRY(cudaEventRecord(ev_b, stream));
            oprtr::LevelCounters c;
            GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, c, stream));
            if (c.overflow)
                return util::GRError(cudaErrorInvalidConfiguration,
                                     "frontier queue overflow: increase queue_size", __FILE__, __LINE__);
            // -- bookkeeping ----------------------------------------------------------------------------
            const unsigned long long D = c.out_len, E = c.edges;
            enactor_stats.total_queued += F;
            enactor_stats.edges_inspected += E;
            if (!pull) {
                ++enactor_stats.push_levels;
                enactor_stats.algo_bytes += F * (4 + 2 * osz) + 5 * E + 8 * D;
            } else {
                ++enactor_stats.pull_levels;
                enactor_stats.algo_bytes += c.aux0 * (2 * osz) + 5 * E + 4 * D + 3 * (n / 8);
            }
            if (opt.record_levels) {
                LevelRecord r;
                r.level = level; r.pull = kind; r.frontier = F; r.edges = E; r.discovered = D; r.candidates = c.aux0;
                GR_TRY(cudaEventElapsedTime(&r.ms, ev_a, ev_b));
                r.switch_ms = 0;
                levels.push_back(r);
            }
            prev_len = F;
            frontier_attribute.queue_length = D;
            frontier_edges = c.out_degree;
            unexplored = unexplored > frontier_edges ? unexplored - frontier_edges : 0;
            edges_visited += frontier_edges;
            frontier_attribute.queue_index++;
            if (pull) cur_map ^= 1; else frontier_attribute.selector ^= 1;
            ++level;
            ++enactor_stats.iteration;
        }
        return cudaSuccess;
    }

    // ---- idempotent mode: advance emits every unmasked neighbour, filter labels + dedups ------
    // (the reference's two-operator pipeline, bfs_enactor.cuh:267-531, bfs_functor.cuh:58-121)
    cudaError_t EnactIdempotent(Problem *problem, const BFSOptions &opt)
    {
        typename Problem::Base::Graph &g = *problem->graph;
        frontier_attribute.queue_length = 1;
        VertexId level = 0;
        const unsigned osz = (unsigned)sizeof(SizeT);
        edges_visited = problem->src_degree;
        while (frontier_attribute.queue_length > 0) {
            const unsigned long long F = frontier_attribute.queue_length;
            DataSlice ds = problem->MakeDataSlice(level);
            AdvArgs a = PushArgs(problem, level);
            oprtr::advance::MODE mode = (opt.traversal_mode == 2) ? oprtr::advance::TWC_FORWARD
                                      : (opt.traversal_mode == 1 || F > 4096 ||
                                         (level == 0 && problem->src_degree >= opt.lb_min_edges)) ? oprtr::advance::LB
                                      : oprtr::advance::TWC_FORWARD;
            if (opt.record_levels) GR_TRY(cudaEventRecord(ev_a, stream));
            GR_TRY((oprtr::advance::LaunchKernel<IdemAdvancePolicy, Problem, BfsFunctor>(
                mode, dev, ds, a, &lb_scratch, (unsigned long long)g.edges, stream)));
            if (opt.record_levels) GR_TRY(cudaEventRecord(ev_b, stream));
            enactor_stats.launches += (mode == oprtr::advance::LB) ? 3 : 2;
            frontier_attribute.queue_index++;
            frontier_attribute.selector ^= 1;
            // filter reads the advance's output length on the device: no host round trip in between
            oprtr::filter::Args<VertexId, Value> fa;
            fa.in_queue = problem->frontier_queues[frontier_attribute.selector].ptr;
            fa.in_pred_queue = nullptr;
            fa.out_queue = problem->frontier_queues[frontier_attribute.selector ^ 1].ptr;
            fa.in_len = problem->frontier_elements[frontier_attribute.selector];
            fa.d_in_len = &work_progress.Slot(frontier_attribute.queue_index)->out_len;
            fa.out_cap = problem->frontier_elements[frontier_attribute.selector ^ 1];
            fa.slots = work_progress.d_slots.ptr;
            fa.queue_index = frontier_attribute.queue_index;
            fa.iteration = level + 1;
            fa.compact = true;
            // big frontiers go through the streaming instance (filter/stream_kernel.cuh); which one this is is only
            // known on the device, so a level whose input frontier is large pays one fetch of the advance's counters
            // in between (~10 us against levels of hundreds)
            oprtr::LevelCounters ca, cf;
            bool filter_streamed = false, have_ca = false;
            if constexpr (Problem::ENABLE_IDEMPOTENCE && !Problem::MARK_PREDECESSORS) {
                if (opt.stream_min_edges && (F >= 32768 || opt.stream_min_edges <= 2) && oprtr::filter::StreamEligible<Problem>(fa, g.part.log_parts)) {
                    GR_TRY(work_progress.Fetch(frontier_attribute.queue_index, ca, stream));
                    have_ca = true;
                    if (!ca.overflow && ca.out_len >= 4 * opt.stream_min_edges) {
                        fa.in_len = ca.out_len; fa.d_in_len = nullptr;
                        filter_streamed = true;
                        ++enactor_stats.launches;                   // kernel + harvest
                        GR_TRY((oprtr::filter::LaunchStreamKernel<IdemStreamPolicy, Problem, BfsFunctor>(
                            dev, ds, fa, g.View(), &lb_scratch, HotWords<IdemStreamPolicy>(problem), stream)));
                    }
                }
            }
            if (!filter_streamed)
                GR_TRY((oprtr::filter::LaunchKernel<FilterPolicy, Problem, BfsFunctor>(dev.sm_count, ds, fa, stream)));
            if (opt.record_levels) GR_TRY(cudaEventRecord(ev_c, stream));
            // one fetch covers both slots (advance wrote queue_index, filter queue_index + 1)
            if (!have_ca) GR_TRY(work_progress.Fetch(frontier_attribute.queue_index, ca, stream));
            GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, cf, stream));
            if (ca.overflow || cf.overflow)
                return util::GRError(cudaErrorInvalidConfiguration,
                                     "frontier queue overflow: increase queue_size", __FILE__, __LINE__);
            enactor_stats.total_queued += F;
            enactor_stats.edges_inspected += ca.edges;
            enactor_stats.algo_bytes += F * (4 + 2 * osz) + 5 * ca.edges + 4 * ca.out_len     // advance
                                      + 5 * ca.out_len + 8 * cf.out_len;                          // filter
            ++enactor_stats.push_levels;
            if (opt.record_levels) {
                // two records per level: the advance (every neighbour emitted) and the filter that culls and
                // compacts them (kind 4: frontier = items read, discovered = items kept)
                LevelRecord ra, rf;
                ra.level = level; ra.pull = (mode == oprtr::advance::LB) ? 2 : 0;
                ra.frontier = F; ra.edges = ca.edges; ra.discovered = ca.out_len; ra.candidates = 0; ra.switch_ms = 0;
                rf.level = level; rf.pull = filter_streamed ? 6 : 4;
                rf.frontier = ca.out_len; rf.edges = 0; rf.discovered = cf.out_len; rf.candidates = 0; rf.switch_ms = 0;
                GR_TRY(cudaEventElapsedTime(&ra.ms, ev_a, ev_b));
                GR_TRY(cudaEventElapsedTime(&rf.ms, ev_b, ev_c));
                levels.push_back(ra); levels.push_back(rf);
            }
            frontier_attribute.queue_length = cf.out_len;
            frontier_attribute.queue_index++;
            frontier_attribute.selector ^= 1;
            ++level;
            ++enactor_stats.iteration;
        }
        return cudaSuccess;
    }
};

}  // namespace bfs
}  // namespace app
}  // namespace gunrock
