// mgbfs_enactor.cuh -- per-level pieces of a 1D vertex-partitioned BFS, one device per process.
//
// The reference declares multi-GPU state (gunrock/app/problem_base.cuh:186-210: owner = v % P,
// local row = v / P) but never implements it (:336-338 "TODO").  This enactor supplies the
// device side of that missing path; the host loop and the collectives live in the harness
// (gunrockinst_b200/partitioned.py: torch.distributed / NCCL over NVLink), which calls the
// steps below on the graph's stream between its collectives.
//
// Every device holds the rows it owns (global neighbour ids) and three GLOBAL 1-bit-per-vertex
// maps laid out owner-major, so that a device's bits are one contiguous, word-aligned segment:
//     visited        own segment authoritative, other segments are cull hints
//     frontier map   x2, own segment written by this device, all-gathered before a pull level
//     next map       discoveries of a push level, whoever owns them
// A level is
//   push:  Push()   advance over the local queue; every unvisited neighbour sets its bit in
//                   the next map (no queue output, no labels yet)
//          --- all-to-all of the next map's segments (harness) ---
//          Absorb() OR of the received segments & ~visited -> labels, own frontier segment,
//                   local queue, counters
//   pull:  --- all-gather of the frontier map (harness) ---
//          Pull()   bottom-up sweep of the owned rows against the global frontier map
// Counters of a level (oprtr::LevelCounters, 64 bytes) sit at a fixed device address that the
// harness all-gathers, so one collective + one host read per level carries the frontier
// length, its degree sum and the edges inspected of every device.
#pragma once
#include <atomic>
#include <chrono>
#include <vector>
#include "bfs_enactor.cuh"
#include "mgbfs_p2p.cuh"
#include "mgbfs_persistent.cuh"

namespace gunrock {
namespace app {
namespace bfs {

// push-level functor: mark the neighbour in the dense discovery map
template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct MgPushFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        const unsigned long long pos = problem->BitPos(d_id);
        const unsigned bit = 1u << (pos & 31);
        unsigned *word = problem->d_next_map + (pos >> 5);
        if (!(util::LoadCA(word) & bit)) atomicOr(word, bit);
        return false;                                   // nothing is enqueued before the exchange
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0) {}
    static __device__ __forceinline__ bool CondFilter(VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) {}
};

// OR the `parts` received copies of this device's segment, keep what is new, label it, publish
// it as this device's segment of the frontier map and enqueue it (CTA-aggregated)
template <typename VertexId, typename SizeT, typename DataSlice, int THREADS>
__global__ void __launch_bounds__(THREADS)
AbsorbDenseKernel(unsigned *recv, int parts, bool consume, unsigned *visited, unsigned *frontier_out, VertexId *queue,
                  unsigned long long cap, const SizeT *row_offsets, oprtr::LevelCounters *slot, DataSlice ds)
{
    __shared__ unsigned s_warp[THREADS / 32 + 1];
    __shared__ unsigned long long s_red[THREADS / 32 + 1];
    __shared__ unsigned long long s_base;
    const long long words = (ds.local_nodes + 31) >> 5;
    const long long seg = ((long long)ds.part.rank * ds.local_nodes) >> 5;
    const long long rounds = (words + (long long)gridDim.x * THREADS - 1) / ((long long)gridDim.x * THREADS);
    unsigned long long my_degree = 0;
    for (long long r = 0; r < rounds; ++r) {
        const long long w = (r * gridDim.x + blockIdx.x) * THREADS + threadIdx.x;
        unsigned bits = 0;
        if (w < words) {
            unsigned acc = 0;
            for (int p = 0; p < parts; ++p) {
                const unsigned x = util::LoadStream(recv + (long long)p * words + w);
                // peer-memory exchange: senders only store the lines that hold discoveries, so a
                // consumed word is cleared here for the next level
                if (consume && x) recv[(long long)p * words + w] = 0;
                acc |= x;
            }
            const unsigned vis = visited[seg + w];
            bits = acc & ~vis;
            frontier_out[seg + w] = bits;
            if (bits) visited[seg + w] = vis | bits;
        }
        unsigned total;
        unsigned excl = util::BlockExclusiveSum<THREADS>((unsigned)__popc(bits), s_warp, total);
        if (total == 0) continue;                       // uniform
        if (threadIdx.x == 0) s_base = atomicAdd(&slot->out_len, (unsigned long long)total);
        __syncthreads();
        unsigned long long at = s_base + excl;
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            const long long row = (w << 5) + b;
            util::StoreCG(ds.d_labels + row, (VertexId)(ds.level + 1));
            my_degree += (unsigned long long)(row_offsets[row + 1] - row_offsets[row]);
            if (at < cap) queue[at] = (VertexId)ds.part.Global(row); else slot->overflow = 1;
            ++at;
        }
        __syncthreads();
    }
    oprtr::advance::BlockAccumulate<THREADS>(my_degree, s_red, &slot->out_degree);
}

// visited |= map over the whole global bitmap (refreshes the cull hints after an all-gather)
static __global__ void MergeVisitedKernel(unsigned *visited, const unsigned *map, unsigned long long words)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long w = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += stride) {
        const unsigned m = map[w];
        if (m) visited[w] |= m;
    }
}

template <typename Problem>
class MGBFSEnactor : public EnactorBase {
  public:
    typedef typename Problem::VertexId VertexId;
    typedef typename Problem::SizeT    SizeT;
    typedef typename Problem::Value    Value;
    typedef typename Problem::DataSlice DataSlice;
    typedef MgPushFunctor<VertexId, SizeT, Value, Problem>     PushFunctor;
    typedef ReverseBFSFunctor<VertexId, SizeT, Value, Problem> PullFunctor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, true, false>  AdvancePolicy;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, GRB_PULL_CTAS, false, false> PullPolicy;

    oprtr::advance::LbScratch<SizeT> lb_scratch;
    int in_sel = 0;                      // frontier_queues[in_sel] is the local frontier
    bool set_up = false;

    oprtr::LevelCounters *Counters() { return work_progress.Slot(1); }

    cudaError_t Setup(Problem *problem, cudaStream_t s)
    {
        GR_TRY(SetupBase(problem->graph->device, s));
        GR_TRY(lb_scratch.Allocate((size_t)problem->graph->local_nodes + 1, (size_t)problem->graph->edges + 1,
                                   AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        GR_TRY(problem->PreparePull());
        GR_TRY(problem->PrepareExchange());
        set_up = true;
        return cudaSuccess;
    }
    cudaError_t Reset(Problem *problem)
    {
        in_sel = 0;
        GR_TRY(ResetBase());
        GR_TRY(problem->frontier_map[0].Zero(stream));
        GR_TRY(problem->frontier_map[1].Zero(stream));
        GR_TRY(problem->next_map.Zero(stream));
        return problem->recv_map.Zero(stream);
    }
    cudaError_t ZeroCounters() { return cudaMemsetAsync(Counters(), 0, sizeof(oprtr::LevelCounters), stream); }

    // push level, part 1: expand the local queue into the dense discovery map
    cudaError_t Push(Problem *problem, VertexId level, unsigned long long queue_len, unsigned long long edges_hint,
                     bool clear_next = true)
    {
        GR_TRY(ZeroCounters());
        if (clear_next) GR_TRY(problem->next_map.Zero(stream));
        if (queue_len == 0) return cudaSuccess;
        DataSlice ds = problem->MakeDataSlice(level);
        oprtr::advance::Args<VertexId, SizeT> a;
        a.g = problem->graph->View();
        a.in_queue = problem->frontier_queues[in_sel].ptr; a.out_queue = nullptr;
        a.in_len = queue_len; a.d_in_len = nullptr; a.out_cap = 0;
        a.slots = work_progress.d_slots.ptr; a.queue_index = 0; a.iteration = level;
        a.output = oprtr::advance::OUT_NONE;
        oprtr::advance::MODE mode = edges_hint >= 4096 ? oprtr::advance::LB : oprtr::advance::TWC_FORWARD;
        enactor_stats.launches += (mode == oprtr::advance::LB) ? 2 : 1;
        return oprtr::advance::LaunchKernel<AdvancePolicy, Problem, PushFunctor>(mode, dev, ds, a, &lb_scratch,
                                                                                 edges_hint, stream);
    }
    // push level, part 2 (after the all-to-all of the next map into recv_map)
    cudaError_t Absorb(Problem *problem, VertexId level, int cur_out, bool consume = false)
    {
        DataSlice ds = problem->MakeDataSlice(level);
        const unsigned long long words = (unsigned long long)((problem->graph->local_nodes + 31) >> 5);
        AbsorbDenseKernel<VertexId, SizeT, DataSlice, 256><<<util::FillGrid(words), 256, 0, stream>>>(
            problem->recv_map.ptr, problem->graph->part.Parts(), consume, problem->visited_mask.ptr,
            problem->frontier_map[cur_out].ptr, problem->frontier_queues[in_sel ^ 1].ptr,
            problem->frontier_elements[in_sel ^ 1], problem->graph->row_offsets.ptr, Counters(), ds);
        GR_TRY(cudaGetLastError());
        in_sel ^= 1;
        ++enactor_stats.launches;
        return cudaSuccess;
    }
    // pull level: reads the (all-gathered) frontier map `cur`, writes this device's segment of map cur^1
    cudaError_t Pull(Problem *problem, VertexId level, int cur, const PeerTable *peers = nullptr)
    {
        GR_TRY(ZeroCounters());
        DataSlice ds = problem->MakeDataSlice(level);
        typename Problem::Base::Graph &g = *problem->graph;
        oprtr::advance::pull::PullArgs<VertexId, SizeT> pa;
        pa.g = g.InverseView();
        pa.frontier_in  = problem->frontier_map[cur].ptr;
        pa.frontier_out = problem->frontier_map[cur ^ 1].ptr;
        pa.visited      = problem->visited_mask.ptr;
        pa.candidates   = problem->candidates.ptr;
        pa.head         = problem->head.ptr;
        pa.tickets      = nullptr;
        pa.fwd_row_offsets = g.symmetric ? nullptr : g.row_offsets.ptr;
        pa.out_queue = nullptr; pa.out_cap = 0;
        pa.peers = 0;
        if (peers)
            for (int p = 0; p < peers->parts; ++p)
                if (p != peers->rank) pa.peer_out[pa.peers++] = peers->map[cur ^ 1][p];
        pa.slots = work_progress.d_slots.ptr; pa.queue_index = 0; pa.iteration = level;
        ++enactor_stats.launches;
        return oprtr::advance::LaunchPullKernel<PullPolicy, Problem, PullFunctor>(dev, ds, pa, stream);
    }
    cudaError_t MergeVisited(Problem *problem, int cur)
    {
        const unsigned long long words = (unsigned long long)((problem->graph->nodes + 31) >> 5);
        MergeVisitedKernel<<<util::FillGrid(words), 256, 0, stream>>>(problem->visited_mask.ptr,
                                                                      problem->frontier_map[cur].ptr, words);
        ++enactor_stats.launches;
        return util::GRError(cudaGetLastError(), "MergeVisitedKernel failed", __FILE__, __LINE__);
    }
    // local queue -> this device's segment of map `cur` (the whole map is cleared first)
    cudaError_t QueueToMap(Problem *problem, int cur, unsigned long long queue_len)
    {
        GR_TRY(problem->frontier_map[cur].Zero(stream));
        if (queue_len == 0) return cudaSuccess;
        DataSlice ds = problem->MakeDataSlice(0);
        QueueToBitmapKernel<VertexId, DataSlice><<<util::FillGrid(queue_len), 256, 0, stream>>>(
            problem->frontier_queues[in_sel].ptr, queue_len, problem->frontier_map[cur].ptr, ds);
        ++enactor_stats.launches;
        return util::GRError(cudaGetLastError(), "QueueToBitmapKernel failed", __FILE__, __LINE__);
    }
    // this device's segment of map `cur` -> local queue (length known to the caller from the counters)
    cudaError_t MapToQueue(Problem *problem, int cur)
    {
        DataSlice ds = problem->MakeDataSlice(0);
        oprtr::LevelCounters *scratch = work_progress.Slot(3);
        GR_TRY(cudaMemsetAsync(scratch, 0, sizeof(oprtr::LevelCounters), stream));
        const unsigned long long words = (unsigned long long)((problem->graph->local_nodes + 31) >> 5);
        BitmapToQueueKernel<VertexId, DataSlice, 256><<<util::FillGrid(words), 256, 0, stream>>>(
            problem->frontier_map[cur].ptr, problem->frontier_queues[in_sel].ptr,
            problem->frontier_elements[in_sel], scratch, ds);
        ++enactor_stats.launches;
        return util::GRError(cudaGetLastError(), "BitmapToQueueKernel failed", __FILE__, __LINE__);
    }

    // ------------------------------------------------------------------------------------------
    // peer-memory exchange (mgbfs_p2p.cuh): the whole search as one native host loop
    // ------------------------------------------------------------------------------------------
    struct LevelTrace {
        int level, kind;                    // kind: 0 push, 1 pull
        unsigned long long frontier, discovered, edges, max_edges, exchange_bytes, candidates;
        float compute_ms, exchange_ms;
    };
    struct SearchStats {
        double elapsed_ms = 0;
        long long iterations = 0, push_levels = 0, pull_levels = 0, launches = 0;
        unsigned long long edges_visited = 0, edges_inspected = 0, exchange_bytes = 0, nodes_visited = 0;
    };

    PeerTable peers;
    util::DeviceArray<Mailbox> mailbox;
    util::DeviceArray<unsigned long long> sent_words;
    util::PinnedArray<oprtr::LevelCounters> h_all;
    util::PinnedArray<unsigned long long> h_status;
    unsigned long long epoch = 0, bar_count = 0, global_edges = 0;
    bool p2p_ready = false;
    // the whole search as one cooperative kernel per device (mgbfs_persistent.cuh)
    static const int kMaxRecords = 128;
    util::DeviceArray<oprtr::LevelCounters> d_all;
    util::DeviceArray<unsigned long long>   d_status;
    util::DeviceArray<MgLevelRecord>        d_records;
    util::DeviceArray<MgPersistentSummary>  d_summary;
    util::PinnedArray<MgLevelRecord>        h_records;
    util::PinnedArray<MgPersistentSummary>  h_mg_summary;
    bool persistent_ok = false;
    int  persistent_grid = 0;
    typedef void (*PersistentFn)(const DataSlice, const MgPersistentParams<VertexId, SizeT>);
    static PersistentFn PersistentEntry() { return MgPersistentKernel<AdvancePolicy, PullPolicy, Problem, PushFunctor, PullFunctor>; }
    static PersistentFn TailEntry() { return MgPersistentKernel<AdvancePolicy, PullPolicy, Problem, PushFunctor, PullFunctor, true>; }
    bool tail_ok = false;
    int  tail_grid = 0;
    static const unsigned long long kTailLimit = 1ull << 16;    // global frontier below which the end of a search runs in one kernel
    std::vector<LevelTrace> trace;
    std::vector<cudaEvent_t> events;

    unsigned long long SegWords(Problem *problem) const { return (unsigned long long)((problem->graph->local_nodes + 31) >> 5); }

    // own entries of the peer table; the remote ones are filled in by the C layer (IPC mappings)
    cudaError_t SetupPeers(Problem *problem)
    {
        GR_TRY(mailbox.Allocate(1));
        GR_TRY(mailbox.Zero(stream));
        GR_TRY(sent_words.Allocate(1));
        GR_TRY(sent_words.Zero(stream));
        GR_TRY(h_all.Allocate(kMaxPeers));
        GR_TRY(h_status.Allocate(1));
        peers = PeerTable();
        peers.parts = problem->graph->part.Parts();
        peers.rank = problem->graph->part.rank;
        if (peers.parts > kMaxPeers) return util::GRError(cudaErrorInvalidValue, "peer exchange supports up to 8 devices", __FILE__, __LINE__);
        const int r = peers.rank;
        peers.mailbox[r] = mailbox.ptr;
        peers.map[0][r] = problem->frontier_map[0].ptr;
        peers.map[1][r] = problem->frontier_map[1].ptr;
        peers.recv[r] = problem->recv_map.ptr;
        epoch = bar_count = 0;
        p2p_ready = (peers.parts == 1);
        GR_TRY(d_all.Allocate(kMaxPeers));
        GR_TRY(d_status.Allocate(1));
        GR_TRY(d_records.Allocate(kMaxRecords));
        GR_TRY(d_summary.Allocate(1));
        GR_TRY(h_records.Allocate(kMaxRecords));
        GR_TRY(h_mg_summary.Allocate(1));
        int coop = 0, per_sm = 0;
        GR_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev.device));
        GR_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, PersistentEntry(), AdvancePolicy::THREADS, 0));
        if (per_sm > kMgPersistentCtasPerSm) per_sm = kMgPersistentCtasPerSm;
        // opt-in (GRB_MG_PERSISTENT=1, bench.py --mg-persistent): measured against the host loop below it saves the host
        // round trip of a level (~15 us) but runs the pull sweep inside a kernel that is register-bound for every phase at
        // once -- slower on one device at scale 24 (1.06 against 0.70 ms per search), see DESIGN.md section 5
        const char *pe = getenv("GRB_MG_PERSISTENT");
        persistent_ok = coop != 0 && per_sm > 0 && pe && pe[0] == '1';
        persistent_grid = per_sm * dev.sm_count;
        // the TAIL instance (the last, short levels of a search in one kernel): opt-in, GRB_MG_TAIL=1.  Measured at 2 devices,
        // scale 27 (profiles/r02_tail_n2.txt): the two last levels take 0.28 ms in the kernel against 0.18 ms driven by the
        // host -- with 2 CTAs per SM the dense scatter / absorb passes over the 16 MB maps are slower than the stand-alone
        // kernels' full grids, and the grid-wide barriers cost what the launches did
        int tail_per_sm = 0;
        GR_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&tail_per_sm, TailEntry(), AdvancePolicy::THREADS, 0));
        if (tail_per_sm > 2) tail_per_sm = 2;               // few CTAs: the grid-wide barriers are what a short level costs
        const char *te = getenv("GRB_MG_TAIL");
        tail_ok = coop != 0 && tail_per_sm > 0 && te && te[0] == '1';
        tail_grid = tail_per_sm * dev.sm_count;
        return cudaStreamSynchronize(stream);
    }

    cudaError_t ExchangeCounters()
    {
        ++epoch;
        *h_status.ptr = 0;
        ExchangeCountersKernel<<<1, 32 * kMaxPeers, 0, stream>>>(peers, Counters(), epoch, h_all.ptr, h_status.ptr,
                                                                 PeerComm::BudgetFromEnv());
        GR_TRY(cudaGetLastError());
        ++enactor_stats.launches;
        // the kernel's last store is the status word in pinned host memory: polling it spares the
        // wake-up latency of a stream synchronisation on every level
        {
            volatile unsigned long long *st = h_status.ptr;
            const auto t0 = std::chrono::steady_clock::now();
            unsigned spins = 0;
            while (*st == 0) {
                if ((++spins & 0xfffu) == 0 &&
                    std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 10.0) break;
            }
            if (*st == 0) GR_TRY(cudaStreamSynchronize(stream));
            std::atomic_thread_fence(std::memory_order_acquire);
        }
        if (*h_status.ptr != epoch)
            return util::GRError(cudaErrorLaunchTimeout, "peer exchange timed out (a peer stopped or the ranks diverged)", __FILE__, __LINE__);
        return cudaSuccess;
    }
    cudaError_t PeerBarrier()
    {
        if (peers.parts == 1) return cudaSuccess;
        ++bar_count;
        PeerBarrierKernel<<<1, 32 * kMaxPeers, 0, stream>>>(peers, bar_count, PeerComm::BudgetFromEnv());
        ++enactor_stats.launches;
        return cudaGetLastError();
    }
    cudaEvent_t Mark(size_t &used)
    {
        if (used == events.size()) { cudaEvent_t e; cudaEventCreate(&e); events.push_back(e); }
        cudaEvent_t e = events[used++];
        cudaEventRecord(e, stream);
        return e;
    }
    void DestroyEvents() { for (cudaEvent_t e : events) cudaEventDestroy(e); events.clear(); }

    void FillPersistentParams(Problem *problem, bool dopt, double alpha, double beta, bool record, MgPersistentParams<VertexId, SizeT> &P)
    {
        typename Problem::Base::Graph &g = *problem->graph;
        P.g = g.View(); P.ginv = g.InverseView();
        P.fwd_row_offsets = g.symmetric ? nullptr : g.row_offsets.ptr;
        for (int i = 0; i < 2; ++i) {
            P.queues[i] = problem->frontier_queues[i].ptr; P.queue_cap[i] = problem->frontier_elements[i];
            P.frontier_map[i] = problem->frontier_map[i].ptr;
        }
        P.visited = problem->visited_mask.ptr; P.next_map = problem->next_map.ptr; P.recv_map = problem->recv_map.ptr;
        P.candidates = problem->candidates.ptr; P.head = problem->head.ptr;
        P.scratch.scanned = lb_scratch.scanned.ptr; P.scratch.row_start = lb_scratch.row_start.ptr;
        P.scratch.tile_desc = lb_scratch.tile_desc.ptr; P.scratch.part_start = lb_scratch.part_start.ptr;
        P.scratch.ticket = lb_scratch.ticket.ptr; P.scratch.part_cap = lb_scratch.part_start.cap;
        P.scratch.epoch = lb_scratch.epoch;
        P.slots = work_progress.d_slots.ptr;
        P.peers = peers;
        P.all = d_all.ptr; P.status = d_status.ptr; P.sent_words = sent_words.ptr;
        P.records = d_records.ptr; P.max_records = record ? kMaxRecords : 0;
        P.summary = d_summary.ptr;
        P.mailbox_epoch = epoch; P.barrier_count = bar_count; P.budget_ns = PeerComm::BudgetFromEnv();
        P.global_edges = global_edges; P.seg_words = SegWords(problem);
        P.desc_words = (unsigned long long)lb_scratch.tile_desc.size;
        P.dopt = dopt ? 1 : 0; P.in_sel = in_sel;
        P.alpha = alpha; P.beta = beta; P.lb_min_edges = 4096;
    }

    // The end of a search in one kernel (MgPersistentKernel<..., TAIL>): called from the host loop with its state after a
    // counter exchange; on return `st` holds the state after the last exchange the kernel made (st.F == 0: search over).
    // -> cudaErrorCooperativeLaunchTooLarge if the launch is refused (the caller goes on level by level).
    cudaError_t RunTail(Problem *problem, bool dopt, double alpha, double beta, bool record, VertexId level_in, MgTailState &st,
                        MgPersistentSummary &sum_out)
    {
        GR_TRY(d_status.Zero(stream));
        MgPersistentParams<VertexId, SizeT> P;
        FillPersistentParams(problem, dopt, alpha, beta, record, P);
        P.tail = st;
        DataSlice ds = problem->MakeDataSlice(level_in);
        void *kargs[] = {(void *)&ds, (void *)&P};
        cudaError_t le = cudaLaunchCooperativeKernel((const void *)TailEntry(), dim3((unsigned)tail_grid),
                                                     dim3(AdvancePolicy::THREADS), kargs, 0, stream);
        if (le != cudaSuccess) {
            (void)cudaGetLastError();
            tail_ok = false;
            return cudaErrorCooperativeLaunchTooLarge;
        }
        ++enactor_stats.launches;
        GR_TRY(cudaMemcpyAsync(h_mg_summary.ptr, d_summary.ptr, sizeof(MgPersistentSummary), cudaMemcpyDeviceToHost, stream));
        if (record)
            GR_TRY(cudaMemcpyAsync(h_records.ptr, d_records.ptr, sizeof(MgLevelRecord) * kMaxRecords, cudaMemcpyDeviceToHost, stream));
        GR_TRY(cudaStreamSynchronize(stream));
        sum_out = *h_mg_summary.ptr;
        epoch = sum_out.mailbox_epoch; bar_count = sum_out.barrier_count;
        in_sel = (int)sum_out.in_sel;
        if (sum_out.error == 1)
            return util::GRError(cudaErrorLaunchTimeout, "peer exchange timed out (a peer stopped or the ranks diverged)", __FILE__, __LINE__);
        if (sum_out.error)
            return util::GRError(cudaErrorInvalidConfiguration, "frontier queue overflow on a partitioned level", __FILE__, __LINE__);
        st.F = sum_out.F; st.frontier_edges = sum_out.frontier_edges; st.local_q = sum_out.local_q; st.local_hint = sum_out.local_hint;
        st.prev_len = sum_out.prev_len; st.unexplored = sum_out.unexplored; st.level = (int)sum_out.level; st.cur = (int)sum_out.cur;
        st.queue_valid = (int)sum_out.queue_valid; st.gathered = (int)sum_out.gathered;
        return cudaSuccess;
    }

    // the search as one cooperative kernel (the reset and the seeding of the counters have been issued on the stream)
    cudaError_t RunPersistent(Problem *problem, bool dopt, double alpha, double beta, bool record, cudaEvent_t ev_start,
                              size_t &used, SearchStats *out)
    {
        GR_TRY(d_status.Zero(stream));
        MgPersistentParams<VertexId, SizeT> P;
        FillPersistentParams(problem, dopt, alpha, beta, record, P);
        P.in_sel = 0;
        DataSlice ds = problem->MakeDataSlice(0);
        void *kargs[] = {(void *)&ds, (void *)&P};
        cudaError_t le = getenv("GRB_FAIL_COOP_LAUNCH") ? cudaErrorCooperativeLaunchTooLarge
                         : cudaLaunchCooperativeKernel((const void *)PersistentEntry(), dim3((unsigned)persistent_grid),
                                                       dim3(AdvancePolicy::THREADS), kargs, 0, stream);
        if (le != cudaSuccess) {
            (void)cudaGetLastError();
            persistent_ok = false;
            fprintf(stderr, "libgunrock (B200 build): cooperative launch of the partitioned BFS kernel failed (%s); "
                            "using the host-driven level loop\n", cudaGetErrorString(le));
            return cudaErrorCooperativeLaunchTooLarge;
        }
        ++enactor_stats.launches;
        GR_TRY(cudaMemcpyAsync(h_mg_summary.ptr, d_summary.ptr, sizeof(MgPersistentSummary), cudaMemcpyDeviceToHost, stream));
        if (record)
            GR_TRY(cudaMemcpyAsync(h_records.ptr, d_records.ptr, sizeof(MgLevelRecord) * kMaxRecords, cudaMemcpyDeviceToHost, stream));
        cudaEvent_t ev_end = Mark(used);
        GR_TRY(cudaEventSynchronize(ev_end));
        const MgPersistentSummary &sum = *h_mg_summary.ptr;
        epoch = sum.mailbox_epoch; bar_count = sum.barrier_count;
        lb_scratch.epoch = (unsigned)sum.lb_epoch_end;
        in_sel = (int)sum.in_sel;
        if (sum.error == 1)
            return util::GRError(cudaErrorLaunchTimeout, "peer exchange timed out (a peer stopped or the ranks diverged)", __FILE__, __LINE__);
        if (sum.error)
            return util::GRError(cudaErrorInvalidConfiguration, "frontier queue overflow on a partitioned level", __FILE__, __LINE__);
        if (record) {
            unsigned long long sent_before = 0;
            const unsigned long long gather = 4ull * SegWords(problem) * (unsigned long long)(peers.parts - 1);
            bool was_pull = false;
            for (unsigned long long i = 0; i < sum.levels && i < (unsigned long long)kMaxRecords; ++i) {
                const MgLevelRecord &r = h_records.ptr[i];
                LevelTrace t;
                t.level = r.level; t.kind = r.kind; t.frontier = r.frontier; t.discovered = r.discovered; t.edges = r.edges;
                t.max_edges = r.max_edges; t.candidates = r.candidates;
                t.exchange_bytes = r.kind ? gather * (was_pull ? 1ull : 2ull) : 4ull * (r.sent_words - sent_before);
                if (peers.parts == 1) t.exchange_bytes = 0;
                sent_before = r.sent_words;
                was_pull = r.kind != 0;
                t.compute_ms = (float)((double)r.ns_compute * 1e-6); t.exchange_ms = (float)((double)r.ns_exchange * 1e-6);
                trace.push_back(t);
            }
        }
        float ms = 0;
        cudaEventElapsedTime(&ms, ev_start, ev_end);
        if (out) {
            out->elapsed_ms = ms; out->iterations = (long long)sum.levels; out->push_levels = (long long)sum.push_levels;
            out->pull_levels = (long long)sum.pull_levels; out->launches = enactor_stats.launches;
            out->edges_visited = sum.edges_visited; out->edges_inspected = sum.edges_inspected;
            out->exchange_bytes = peers.parts > 1 ? sum.gather_bytes : 0; out->nodes_visited = sum.nodes_visited;
        }
        return cudaSuccess;
    }

    // One direction-optimising search; every rank calls it with the same arguments.  The direction
    // rule is the one of BFSEnactor (Beamer's alpha / beta on global frontier figures).
    cudaError_t RunPeer(Problem *problem, VertexId src, bool dopt, double alpha, double beta, bool record, SearchStats *out)
    {
        if (!p2p_ready) return util::GRError(cudaErrorNotReady, "peer mappings have not been imported", __FILE__, __LINE__);
        typename Problem::Base::Graph &g = *problem->graph;
        if (dopt && !g.HasInverse()) {          // a directed slice: the pull direction needs in-edges this device does not hold
            static bool told = false;
            if (!told) fprintf(stderr, "libgunrock (B200 build): partitioned BFS on a directed graph runs top-down only "
                                       "(no per-device transpose)\n");
            told = true;
            dopt = false;
        }
        const int P = peers.parts;
        const unsigned long long seg_words = SegWords(problem);
        const unsigned long long n = (unsigned long long)g.nodes;
        size_t used = 0;
        trace.clear();
        cudaEvent_t ev_start = Mark(used);

        // ---- reset: no host round trip; the first counter exchange is the barrier behind it
        GR_TRY(problem->Reset(-1, VERTEX_FRONTIERS, 1.0, true));
        problem->src = src;
        in_sel = 0;
        GR_TRY(ResetBase());
        if (g.local_nodes) GR_TRY(cudaMemsetAsync(problem->labels.ptr, 0xff, sizeof(VertexId) * (size_t)g.local_nodes, stream));
        GR_TRY(problem->visited_mask.Zero(stream));
        GR_TRY(problem->frontier_map[0].Zero(stream));
        GR_TRY(problem->next_map.Zero(stream));
        GR_TRY(problem->recv_map.Zero(stream));
        GR_TRY(sent_words.Zero(stream));
        unsigned long long sent_so_far = 0;
        const bool valid = src >= 0 && (unsigned long long)src < n;
        const bool mine = valid && g.part.Mine(src);
        {
            DataSlice ds = problem->MakeDataSlice(0);
            SeedSourceKernel<VertexId, DataSlice><<<1, 32, 0, stream>>>(ds, src, problem->frontier_queues[0].ptr, problem->frontier_map[0].ptr);
            SeedCountersKernel<SizeT><<<1, 1, 0, stream>>>(Counters(), g.row_offsets.ptr, mine ? (long long)g.part.Local(src) : -1ll);
            GR_TRY(cudaGetLastError());
            enactor_stats.launches += 2;
        }
        if (persistent_ok) {
            cudaError_t pe = RunPersistent(problem, dopt, alpha, beta, record, ev_start, used, out);
            if (pe != cudaErrorCooperativeLaunchTooLarge) return pe;
            // co-residency is not available after all: the host loop below runs this and every later search.  (Every
            // device takes this branch or none: the launch fails for a reason all devices of a box share, and a device
            // that went ahead alone would time out at its first exchange.)
        }
        GR_TRY(ExchangeCounters());

        unsigned long long frontier = 0, frontier_edges = 0;
        for (int p = 0; p < P; ++p) { frontier += h_all.ptr[p].out_len; frontier_edges += h_all.ptr[p].out_degree; }
        unsigned long long local_q = h_all.ptr[peers.rank].out_len, local_hint = h_all.ptr[peers.rank].out_degree;
        unsigned long long unexplored = global_edges > frontier_edges ? global_edges - frontier_edges : 0;
        unsigned long long edges_visited = frontier_edges, nodes_visited = frontier, inspected = 0, exch_bytes = 0, prev_len = 0;
        bool pull = false, gathered = false, queue_valid = true;
        int cur = 0;
        long long level = 0, push_levels = 0, pull_levels = 0;

        while (frontier > 0) {
            if (dopt) {
                if (!pull && (double)frontier_edges > (double)unexplored / alpha) pull = true;
                else if (pull && frontier < prev_len && (double)frontier < (double)n / beta) pull = false;
            }
            // ---- the end of the search: once the frontier is short and shrinking, the remaining levels run in one kernel
            if (tail_ok && !pull && level > 0 && frontier < kTailLimit && frontier < prev_len) {
                MgTailState ts;
                ts.F = frontier; ts.frontier_edges = frontier_edges; ts.local_q = local_q; ts.local_hint = local_hint;
                ts.prev_len = prev_len; ts.unexplored = unexplored; ts.limit = kTailLimit;
                ts.level = (int)level; ts.cur = cur; ts.queue_valid = queue_valid ? 1 : 0; ts.gathered = gathered ? 1 : 0;
                MgPersistentSummary ts_sum;
                cudaError_t te = RunTail(problem, dopt, alpha, beta, record, (VertexId)level, ts, ts_sum);
                if (te == cudaSuccess) {
                    if (record) {
                        for (unsigned long long i = 0; i < ts_sum.levels && i < (unsigned long long)kMaxRecords; ++i) {
                            const MgLevelRecord &r = h_records.ptr[i];
                            LevelTrace t;
                            t.level = r.level; t.kind = r.kind; t.frontier = r.frontier; t.discovered = r.discovered; t.edges = r.edges;
                            t.max_edges = r.max_edges; t.candidates = 0;
                            t.exchange_bytes = P > 1 ? 4ull * (r.sent_words - sent_so_far) : 0ull;
                            sent_so_far = r.sent_words;
                            exch_bytes += t.exchange_bytes;
                            t.compute_ms = (float)((double)r.ns_compute * 1e-6); t.exchange_ms = (float)((double)r.ns_exchange * 1e-6);
                            trace.push_back(t);
                        }
                    }
                    push_levels += (long long)ts_sum.push_levels;
                    inspected += ts_sum.edges_inspected;
                    edges_visited += ts_sum.edges_visited;
                    nodes_visited += ts_sum.nodes_visited;
                    frontier = ts.F; frontier_edges = ts.frontier_edges; local_q = ts.local_q; local_hint = ts.local_hint;
                    prev_len = ts.prev_len; unexplored = ts.unexplored; level = ts.level; cur = ts.cur;
                    queue_valid = ts.queue_valid != 0; gathered = ts.gathered != 0;
                    continue;                       // frontier == 0: done; else the kernel handed the loop back
                }
                if (te != cudaErrorCooperativeLaunchTooLarge) return te;
            }
            cudaEvent_t m0 = 0, m1 = 0, m2 = 0, m3 = 0, m4 = 0;
            unsigned long long lvl_bytes = 0;
            if (record) m0 = Mark(used);
            if (pull) {
                if (record) m1 = Mark(used);
                if (!gathered && P > 1) {       // first pull level after push levels: own segment -> every peer
                    ScatterSegmentKernel<<<util::FillGrid(seg_words >> 2), 256, 0, stream>>>(peers, cur, seg_words);
                    GR_TRY(cudaGetLastError());
                    ++enactor_stats.launches;
                    GR_TRY(PeerBarrier());
                    lvl_bytes += 4ull * seg_words * (unsigned long long)(P - 1);
                }
                if (record) m2 = Mark(used);
                GR_TRY(MergeVisited(problem, cur));
                GR_TRY(Pull(problem, (VertexId)level, cur, &peers));      // publishes map cur^1 on every device
                lvl_bytes += 4ull * seg_words * (unsigned long long)(P - 1);
                cur ^= 1;
                queue_valid = false;
                gathered = true;
                ++pull_levels;
            } else {
                if (!queue_valid) { GR_TRY(MapToQueue(problem, cur)); queue_valid = true; }
                GR_TRY(Push(problem, (VertexId)level, local_q, local_hint, false));
                if (record) m1 = Mark(used);
                ScatterNextKernel<<<util::FillGrid(seg_words * (unsigned long long)P), 256, 0, stream>>>(
                    peers, problem->next_map.ptr, seg_words, sent_words.ptr);
                GR_TRY(cudaGetLastError());
                ++enactor_stats.launches;
                GR_TRY(PeerBarrier());
                if (record) m2 = Mark(used);
                GR_TRY(Absorb(problem, (VertexId)level, cur ^ 1, true));
                cur ^= 1;
                gathered = false;
                ++push_levels;
            }
            if (record) m3 = Mark(used);
            GR_TRY(ExchangeCounters());
            if (record) m4 = Mark(used);
            unsigned long long f = 0, fe = 0, ins = 0, max_ins = 0, ovf = 0, cands = 0;
            for (int p = 0; p < P; ++p) {
                const oprtr::LevelCounters &c = h_all.ptr[p];
                f += c.out_len; fe += c.out_degree; ins += c.edges; ovf += c.overflow; cands += c.aux0;
                if (c.edges > max_ins) max_ins = c.edges;
            }
            if (ovf) return util::GRError(cudaErrorInvalidConfiguration, "frontier queue overflow on a partitioned level", __FILE__, __LINE__);
            if (!pull && record) {  // bytes this device actually sent: the lines of the discovery map that held a bit
                unsigned long long sw = 0;
                GR_TRY(cudaMemcpyAsync(&sw, sent_words.ptr, sizeof(sw), cudaMemcpyDeviceToHost, stream));
                GR_TRY(cudaStreamSynchronize(stream));
                lvl_bytes += 4ull * (sw - sent_so_far);
                sent_so_far = sw;
            }
            if (record) {
                LevelTrace t;
                t.level = (int)level; t.kind = pull ? 1 : 0;
                t.frontier = frontier; t.discovered = f; t.edges = ins; t.max_edges = max_ins; t.exchange_bytes = lvl_bytes; t.candidates = pull ? cands : 0;
                float a = 0, b = 0, c = 0, d = 0;
                cudaEventSynchronize(m4);
                cudaEventElapsedTime(&a, m0, m1); cudaEventElapsedTime(&b, m1, m2);
                cudaEventElapsedTime(&c, m2, m3); cudaEventElapsedTime(&d, m3, m4);
                t.compute_ms = a + c; t.exchange_ms = b + d;
                trace.push_back(t);
            }
            prev_len = frontier;
            frontier = f; frontier_edges = fe;
            local_q = h_all.ptr[peers.rank].out_len; local_hint = h_all.ptr[peers.rank].out_degree;
            inspected += ins;
            unexplored = unexplored > frontier_edges ? unexplored - frontier_edges : 0;
            edges_visited += frontier_edges;
            nodes_visited += frontier;
            exch_bytes += lvl_bytes;
            ++level;
        }
        if (!record && push_levels) {   // push payload of the whole search (the traced path counted it per level)
            unsigned long long sw = 0;
            GR_TRY(cudaMemcpyAsync(&sw, sent_words.ptr, sizeof(sw), cudaMemcpyDeviceToHost, stream));
            GR_TRY(cudaStreamSynchronize(stream));
            exch_bytes += 4ull * sw;
        }
        cudaEvent_t ev_end = Mark(used);
        GR_TRY(cudaEventSynchronize(ev_end));
        float ms = 0;
        cudaEventElapsedTime(&ms, ev_start, ev_end);
        if (out) {
            out->elapsed_ms = ms; out->iterations = level; out->push_levels = push_levels; out->pull_levels = pull_levels;
            out->launches = enactor_stats.launches; out->edges_visited = edges_visited; out->edges_inspected = inspected;
            out->exchange_bytes = exch_bytes; out->nodes_visited = nodes_visited;
        }
        return cudaSuccess;
    }
};

}  // namespace bfs
}  // namespace app
}  // namespace gunrock
