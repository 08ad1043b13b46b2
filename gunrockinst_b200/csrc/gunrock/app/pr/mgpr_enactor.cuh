// mgpr_enactor.cuh -- 1D vertex-partitioned PageRank, one device per process, rank
// contributions published through peer memory (NVLink / NVSwitch).
//
// No reference counterpart for the partitioned part (gunrock/app/problem_base.cuh:336-338 is a
// TODO).  The arithmetic is the reference's (gunrock/app/pr/pr_functor.cuh:32-175,
// pr_enactor.cuh:247-329,346-487, pr_problem.cuh:407-413) as restated by the single-device
// PREnactor (pr_enactor.cuh in this directory), in fp64:
//   prune   vertices without out-edges leave, their in-neighbours' degrees drop, repeated until
//           nothing leaves;
//   sweep   next[v] = delta * sum_{u -> v, u alive} curr[u] / degree[u] + (1 - delta) * base(v),
//           for alive v; a vertex is "moving" while |next - curr| > threshold; the loop ends when
//           no device has a moving vertex or after max_iter sweeps.
//
// The single-device code pushes along out-edges with one atomicAdd per edge into the
// destination.  Partitioned, a push would be a remote fp64 atomic per cut edge; this enactor
// GATHERS instead: every device keeps a replica of all contributions curr[u] / degree[u]
// (8 bytes per global vertex, owner-major index, two generations) and sums, for the rows it
// owns, over their in-edges -- which on a symmetric graph are the rows it already holds.  The
// only traffic is the publication of the new contributions: the kernel that finishes a sweep
// stores each owned vertex's value into every device's replica (coalesced 8-byte stores over
// NVLink, n / P * 8 bytes per peer and sweep), so the exchange is fused into the compute kernel
// and no collective runs.  The pruning phase works the same way with a replicated 1-bit
// "out-degree just reached zero" map.  Per sweep and device:
//     advance   LB operator over the owned alive rows, functor adds contrib[neighbour] into the
//               row's accumulator (warp-aggregated when the whole warp works on one row)
//     finish    per owned row: damping, moving test, new contribution -> all replicas
//     exchange  64-byte counter blocks of all devices (also the barrier behind the stores)
// Needs a symmetric graph (in-edges == out-edges), like the partitioned BFS.
#pragma once
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../enactor_base.cuh"
#include "../bfs/mgbfs_p2p.cuh"
#include "../../util/memset_kernel.cuh"
#include "pr_row_sweep.cuh"

namespace gunrock {
namespace app {
namespace pr {

struct PrPeerTable {
    double   *contrib[bfs::kMaxPeers] = {};     // [2][parts * local_nodes] on every device
    unsigned *zero[bfs::kMaxPeers]    = {};     // [2][parts * local_nodes / 32] on every device
    int       parts = 1, rank = 0;
};

template <typename _SizeT>
struct MgPrProblem {
    typedef int     VertexId;
    typedef _SizeT  SizeT;
    typedef double  Value;
    static const bool MARK_PREDECESSORS  = true;    // the operators hand the real source vertex to the functor
    static const bool ENABLE_IDEMPOTENCE = false;
    struct DataSlice {
        const double   *contrib_in;     // replica read by this sweep
        double         *acc;            // per owned row: sum of the in-neighbours' contributions
        const unsigned *zero_in;        // replica read by this pruning sweep
        int            *deg;            // per owned row: reduced out-degree, < 0 once pruned
        long long       nodes, local_nodes;
        graph::Partition part;
    };
};

// position of global vertex v in the owner-major replicas
template <typename DataSlice>
__device__ __forceinline__ unsigned long long ReplicaPos(const DataSlice *p, int v)
{
    return (unsigned long long)p->part.Owner(v) * (unsigned long long)p->local_nodes +
           (unsigned long long)p->part.Local(v);
}

// edge (s, d) of an owned row s: d -> s exists as well (symmetric graph), so d's contribution
// flows into s.  When all 32 lanes hold edges of the same row (every row of >= 32 entries, for
// most of its length) the warp adds up first and issues one atomic.
template <typename VertexId, typename SizeT, typename ProblemData>
struct MgPrGatherFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *p, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        double c = __ldg(p->contrib_in + ReplicaPos(p, d_id));
        const unsigned act = __activemask();
        if (act == util::kFullMask) {
            const VertexId s0 = __shfl_sync(util::kFullMask, s_id, 0);
            if (__all_sync(util::kFullMask, s_id == s0)) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(util::kFullMask, c, o);
                if (util::LaneId() == 0 && c != 0.0) atomicAdd(p->acc + p->part.Local(s_id), c);
                return false;
            }
        }
        if (c != 0.0) atomicAdd(p->acc + p->part.Local(s_id), c);
        return false;
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *p, SizeT e_id = 0, SizeT e_id_in = 0) {}
    static __device__ __forceinline__ bool CondFilter(VertexId node, const DataSlice *p, double v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(VertexId node, const DataSlice *p, double v = 0, SizeT nid = 0) {}
};

// pruning: an out-neighbour whose degree just reached zero takes one off this row's degree
// (gunrock/app/pr/pr_functor.cuh:113-175)
template <typename VertexId, typename SizeT, typename ProblemData>
struct MgPrPruneFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *p, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        if (util::BitTest(p->zero_in, ReplicaPos(p, d_id))) atomicAdd(p->deg + p->part.Local(s_id), -1);
        return false;
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *p, SizeT e_id = 0, SizeT e_id_in = 0) {}
    static __device__ __forceinline__ bool CondFilter(VertexId node, const DataSlice *p, double v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(VertexId node, const DataSlice *p, double v = 0, SizeT nid = 0) {}
};

// one word of a replicated bitmap, stored on every device
__device__ __forceinline__ void PublishWord(const PrPeerTable &t, unsigned long long half_words, int gen,
                                            unsigned long long word, unsigned value)
{
    for (int p = 0; p < t.parts; ++p) t.zero[p][(unsigned long long)gen * half_words + word] = value;
}

// degrees, ranks, accumulators of the owned rows; identity queue; generation-0 "degree is zero" map
template <typename SizeT>
__global__ void __launch_bounds__(256)
MgPrInitKernel(const SizeT *row_offsets, long long local_nodes, graph::Partition part, PrPeerTable peers,
               int *deg, double *rank, double *acc, int *queue, double init, oprtr::LevelCounters *c)
{
    __shared__ unsigned long long s_red[256 / 32 + 1];
    const unsigned long long half_words = ((unsigned long long)local_nodes * peers.parts + 31) >> 5;
    const unsigned long long seg = ((unsigned long long)part.rank * (unsigned long long)local_nodes) >> 5;
    const long long rounded = (local_nodes + 31) & ~31ll;
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned long long zeros = 0;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rounded; r += stride) {
        bool z = false;
        if (r < local_nodes) {
            const int d = (int)(row_offsets[r + 1] - row_offsets[r]);
            deg[r] = d; rank[r] = init; acc[r] = 0.0;
            queue[r] = (int)part.Global(r);
            z = d == 0;
        }
        const unsigned m = __ballot_sync(util::kFullMask, z);
        if (util::LaneId() == 0) { PublishWord(peers, half_words, 0, seg + (unsigned long long)(r >> 5), m); zeros += (unsigned)__popc(m); }
    }
    oprtr::advance::BlockAccumulate<256>(zeros, s_red, &c->aux1);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&c->out_len, (unsigned long long)local_nodes);
}

// end of a pruning sweep: rows whose degree was zero when the sweep began leave (degree -1), the
// others stay with what the advance left them; the next generation of the map marks the rows
// that have just reached zero.  out_len = rows that stay, aux1 = bits set in the new map.
static __global__ void __launch_bounds__(256)
MgPrPruneFinishKernel(long long local_nodes, graph::Partition part, PrPeerTable peers, int gen_in, int *deg,
                      oprtr::LevelCounters *c)
{
    __shared__ unsigned long long s_red[256 / 32 + 1];
    const unsigned long long half_words = ((unsigned long long)local_nodes * peers.parts + 31) >> 5;
    const unsigned long long seg = ((unsigned long long)part.rank * (unsigned long long)local_nodes) >> 5;
    const unsigned *zero_in = peers.zero[peers.rank] + (unsigned long long)gen_in * half_words;
    const long long rounded = (local_nodes + 31) & ~31ll;
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned long long kept = 0, zeros = 0;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < rounded; r += stride) {
        const unsigned zw = zero_in[seg + (unsigned long long)(r >> 5)];      // same word for the whole warp
        bool z = false;
        if (r < local_nodes) {
            int d = deg[r];
            if (d >= 0 && ((zw >> (r & 31)) & 1u)) { d = -1; deg[r] = -1; }
            if (d >= 0) { ++kept; z = d == 0; }
        }
        const unsigned m = __ballot_sync(util::kFullMask, z);
        if (util::LaneId() == 0) { PublishWord(peers, half_words, gen_in ^ 1, seg + (unsigned long long)(r >> 5), m); zeros += (unsigned)__popc(m); }
    }
    oprtr::advance::BlockAccumulate<256>(kept, s_red, &c->out_len);
    oprtr::advance::BlockAccumulate<256>(zeros, s_red, &c->aux1);
}

// after the pruning: queue of the surviving rows (global ids), first contributions
// (curr / degree, 0 for pruned rows -- these in both generations) into generation 0 of every replica; a pruned row's rank is
// what the single-device loop leaves it with after its first sweep, 0
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
MgPrStartKernel(long long local_nodes, graph::Partition part, PrPeerTable peers, const int *deg, double *rank,
                int *queue, oprtr::LevelCounters *c)
{
    __shared__ unsigned s_warp[THREADS / 32 + 1];
    __shared__ unsigned long long s_base;
    const unsigned long long first = (unsigned long long)part.rank * (unsigned long long)local_nodes;
    const unsigned long long half = (unsigned long long)local_nodes * (unsigned long long)peers.parts;
    const long long rounds = (local_nodes + (long long)gridDim.x * THREADS - 1) / ((long long)gridDim.x * THREADS);
    for (long long k = 0; k < rounds; ++k) {
        const long long r = (k * gridDim.x + blockIdx.x) * THREADS + threadIdx.x;
        bool alive = false;
        if (r < local_nodes) {
            const int d = deg[r];
            alive = d > 0;
            double cv = 0.0;
            if (alive) cv = rank[r] / (double)d; else rank[r] = 0.0;
            for (int p = 0; p < peers.parts; ++p) peers.contrib[p][first + (unsigned long long)r] = cv;
            // (a sweep that only visits the alive rows -- pr_row_sweep.cuh -- reads both generations of a pruned row)
            if (!alive) for (int p = 0; p < peers.parts; ++p) peers.contrib[p][half + first + (unsigned long long)r] = 0.0;
        }
        unsigned total;
        const unsigned excl = util::BlockExclusiveSum<THREADS>(alive ? 1u : 0u, s_warp, total);
        if (total == 0) continue;                       // uniform
        if (threadIdx.x == 0) s_base = atomicAdd(&c->out_len, (unsigned long long)total);
        __syncthreads();
        if (alive) queue[s_base + excl] = (int)part.Global(r);
        __syncthreads();
    }
}

// end of a sweep, per owned alive row: damping and base term, moving test, the new contribution
// into generation gen_out of every replica, accumulator back to zero
// (gunrock/app/pr/pr_functor.cuh:73-100)
static __global__ void __launch_bounds__(256)
MgPrFinishKernel(long long local_nodes, graph::Partition part, PrPeerTable peers, int gen_out, const int *deg,
                 double *rank, double *acc, double delta, double threshold, int src, oprtr::LevelCounters *c)
{
    __shared__ unsigned long long s_red[256 / 32 + 1];
    const unsigned long long half = (unsigned long long)local_nodes * (unsigned long long)peers.parts;
    const unsigned long long first = (unsigned long long)gen_out * half + (unsigned long long)part.rank * (unsigned long long)local_nodes;
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned long long moving = 0;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < local_nodes; r += stride) {
        const int d = deg[r];
        double cv = 0.0;
        if (d > 0) {
            const int v = (int)part.Global(r);
            const double base = (src == v || src == -1) ? 1.0 : 0.0;
            const double next = __dadd_rn(__dmul_rn(delta, acc[r]), __dmul_rn(1.0 - delta, base));
            acc[r] = 0.0;
            if (fabs(next - rank[r]) > threshold) ++moving;
            rank[r] = next;
            cv = next / (double)d;
        }
        for (int p = 0; p < peers.parts; ++p) peers.contrib[p][first + (unsigned long long)r] = cv;
    }
    oprtr::advance::BlockAccumulate<256>(moving, s_red, &c->aux0);
}

template <typename _SizeT>
class MGPREnactor : public EnactorBase {
  public:
    typedef _SizeT SizeT;
    typedef int    VertexId;
    typedef MgPrProblem<SizeT> Problem;
    typedef typename Problem::DataSlice DataSlice;
    typedef MgPrGatherFunctor<VertexId, SizeT, Problem> GatherFunctor;
    typedef MgPrPruneFunctor<VertexId, SizeT, Problem>  PruneFunctor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, false> AdvancePolicy;
    typedef graph::DeviceCsr<VertexId, SizeT> Graph;

    struct RunStats {
        double elapsed_ms = 0, sweep_ms = 0;
        long long iterations = 0, prune_sweeps = 0, launches = 0;
        unsigned long long edges = 0, valid_nodes = 0, exchange_bytes = 0;
        bool by_rows = false;           // the sweeps ran as pr_row_sweep.cuh's kernel
    };

    Graph *graph = nullptr;
    util::DeviceArray<double>   contrib, rank, acc;     // contrib: [2][parts * local_nodes] (peer-mapped)
    util::DeviceArray<unsigned> zero;                   // [2][parts * local_nodes / 32]     (peer-mapped)
    util::DeviceArray<int>      deg;
    util::DeviceArray<VertexId> queue;
    oprtr::advance::LbScratch<SizeT> lb_scratch;
    // single device: the rows of the power iteration binned by length (pr_row_sweep.cuh)
    bool row_sweep = false, row_tickets = false;
    oprtr::rows::Bins<SizeT> row_bins;
    util::DeviceArray<unsigned long long> sweep_ticket;
    int ticket_ctas_per_sm = 0;
    bfs::PeerComm comm;
    PrPeerTable peers;
    unsigned long long half = 0, half_words = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;

    ~MGPREnactor()
    {
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        if (ev2) cudaEventDestroy(ev2);
    }
    oprtr::LevelCounters *Counters() { return work_progress.Slot(1); }

    cudaError_t Setup(Graph *g, cudaStream_t s)
    {
        graph = g;
        if (!g->symmetric)
            return util::GRError(cudaErrorInvalidValue, "partitioned PageRank needs a symmetric graph", __FILE__, __LINE__);
        if (g->local_nodes % 32 != 0 && g->part.log_parts != 0)
            return util::GRError(cudaErrorInvalidValue, "partition segments must be word aligned", __FILE__, __LINE__);
        GR_TRY(SetupBase(g->device, s));
        const size_t ln = (size_t)g->local_nodes, P = (size_t)g->part.Parts();
        half = (unsigned long long)ln * P;
        half_words = (half + 31) >> 5;
        GR_TRY(contrib.Allocate(2 * (size_t)half + 1));
        GR_TRY(zero.Allocate(2 * (size_t)half_words + 1));
        GR_TRY(rank.Allocate(ln + 1)); GR_TRY(acc.Allocate(ln + 1));
        GR_TRY(deg.Allocate(ln + 1));  GR_TRY(queue.Allocate(ln + 1));
        GR_TRY(lb_scratch.Allocate(ln + 1, (size_t)g->edges + 1, AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        const char *route = getenv("GRB_PR_SWEEP");
        row_sweep = P == 1 && !(route && strcmp(route, "lb") == 0);
        row_tickets = row_sweep && !(route && strcmp(route, "rows1") == 0);     // rows1: the static grid of PrRowSweepKernel
        if (row_sweep) {
            GR_TRY(row_bins.Allocate(ln, (size_t)g->edges));
            GR_TRY(sweep_ticket.Allocate(1));
            GR_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ticket_ctas_per_sm, PrRowSweepTicketKernel<SizeT>,
                                                                 oprtr::rows::kThreads, 0));
            if (ticket_ctas_per_sm < 1) ticket_ctas_per_sm = 1;
        }
        GR_TRY(comm.Setup((int)P, g->part.rank, s));
        peers = PrPeerTable();
        peers.parts = (int)P; peers.rank = g->part.rank;
        peers.contrib[g->part.rank] = contrib.ptr;
        peers.zero[g->part.rank] = zero.ptr;
        GR_TRY(cudaEventCreate(&ev0)); GR_TRY(cudaEventCreate(&ev1)); GR_TRY(cudaEventCreate(&ev2));
        return cudaStreamSynchronize(s);
    }

    DataSlice MakeDataSlice(int contrib_gen, int zero_gen) const
    {
        DataSlice d;
        d.contrib_in = contrib.ptr + (unsigned long long)contrib_gen * half;
        d.acc = acc.ptr;
        d.zero_in = zero.ptr + (unsigned long long)zero_gen * half_words;
        d.deg = deg.ptr;
        d.nodes = graph->nodes; d.local_nodes = graph->local_nodes; d.part = graph->part;
        return d;
    }

    template <typename Functor>
    cudaError_t Advance(const DataSlice &ds, unsigned long long qlen, long long iteration)
    {
        Graph &g = *graph;
        oprtr::advance::Args<VertexId, SizeT> a;
        a.g = g.View();
        a.in_queue = queue.ptr; a.out_queue = nullptr;
        a.in_len = qlen; a.d_in_len = nullptr; a.out_cap = 0;
        a.slots = work_progress.d_slots.ptr; a.queue_index = 0; a.iteration = iteration;
        a.output = oprtr::advance::OUT_NONE;
        enactor_stats.launches += 2;
        return oprtr::advance::LaunchKernel<AdvancePolicy, Problem, Functor>(
            oprtr::advance::LB, dev, ds, a, &lb_scratch, (unsigned long long)g.edges, stream);
    }

    // the alive rows (queue[0 .. qlen)) into the three bins of oprtr/row_bins.cuh; fills everything of `out` that
    // does not change from sweep to sweep
    cudaError_t BinRows(unsigned long long qlen, double delta, double threshold, VertexId src, PrRowSweepArgs<SizeT> *out)
    {
        Graph &g = *graph;
        GR_TRY(row_bins.Build(g.row_offsets.ptr, queue.ptr, qlen, g.part, stream));
        ++enactor_stats.launches;
        PrRowSweepArgs<SizeT> a;
        a.bins = row_bins.view;
        if (a.bins.n_small + a.bins.n_medium + a.bins.n_big != qlen)       // every alive row has entries
            return util::GRError(cudaErrorUnknown, "PageRank row bins do not add up", __FILE__, __LINE__);
        a.row_offsets = g.row_offsets.ptr; a.col_indices = g.col_indices.ptr;
        a.contrib_in = nullptr; a.contrib_out = nullptr;
        a.deg = deg.ptr; a.rank = rank.ptr; a.acc = acc.ptr;
        a.delta = delta; a.threshold = threshold; a.src = src;
        a.part = g.part; a.local_nodes = g.local_nodes;
        a.counters = Counters();
        *out = a;
        return cudaSuccess;
    }

    // every rank calls it with the same arguments
    cudaError_t Run(VertexId src, double delta, double threshold, long long max_iteration, RunStats *out)
    {
        if (!comm.ready) return util::GRError(cudaErrorNotReady, "peer mappings have not been imported", __FILE__, __LINE__);
        Graph &g = *graph;
        const long long ln = g.local_nodes;
        const int P = comm.table.parts, me = comm.table.rank;
        const unsigned long long row_grid = (unsigned long long)ln + 32;
        RunStats st;
        GR_TRY(cudaEventRecord(ev0, stream));
        GR_TRY(ResetBase());
        GR_TRY(cudaMemsetAsync(Counters(), 0, sizeof(oprtr::LevelCounters), stream));
        MgPrInitKernel<SizeT><<<util::FillGrid(row_grid), 256, 0, stream>>>(
            g.row_offsets.ptr, ln, g.part, peers, deg.ptr, rank.ptr, acc.ptr, queue.ptr, 1.0 - delta, Counters());
        GR_TRY(cudaGetLastError());
        GR_TRY(comm.Exchange(Counters(), stream));      // also the barrier behind every device's generation-0 map
        enactor_stats.launches += 2;

        // ---- (a) prune vertices without out-edges (pr_enactor.cuh:247-329) --------------------------
        unsigned long long qlen = 0, zeros = 0, prev = 0, local_q = (unsigned long long)ln;
        for (int p = 0; p < P; ++p) { qlen += comm.h_all.ptr[p].out_len; zeros += comm.h_all.ptr[p].aux1; }
        int zgen = 0;
        while (prev != qlen) {
            prev = qlen;
            ++st.prune_sweeps;
            if (zeros == 0) break;          // nothing reached zero in the last sweep: this one would remove nothing
            GR_TRY(cudaMemsetAsync(Counters(), 0, sizeof(oprtr::LevelCounters), stream));
            // (the identity queue of the first sweep is kept: rows that have left only go further below zero)
            if (local_q) GR_TRY(Advance<PruneFunctor>(MakeDataSlice(0, zgen), local_q, 0));
            MgPrPruneFinishKernel<<<util::FillGrid(row_grid), 256, 0, stream>>>(ln, g.part, peers, zgen, deg.ptr, Counters());
            GR_TRY(cudaGetLastError());
            GR_TRY(comm.Exchange(Counters(), stream));
            enactor_stats.launches += 2;
            st.exchange_bytes += 4ull * (unsigned long long)((ln + 31) >> 5) * (unsigned long long)(P - 1);
            qlen = 0; zeros = 0;
            for (int p = 0; p < P; ++p) { qlen += comm.h_all.ptr[p].out_len; zeros += comm.h_all.ptr[p].aux1; st.edges += comm.h_all.ptr[p].edges; }
            zgen ^= 1;
        }
        st.valid_nodes = qlen;
        bool swept = false;
        if (qlen > 0 && max_iteration > 0) {
            GR_TRY(cudaMemsetAsync(Counters(), 0, sizeof(oprtr::LevelCounters), stream));
            MgPrStartKernel<256><<<util::FillGrid(row_grid), 256, 0, stream>>>(ln, g.part, peers, deg.ptr, rank.ptr, queue.ptr, Counters());
            GR_TRY(cudaGetLastError());
            GR_TRY(comm.Exchange(Counters(), stream));
            enactor_stats.launches += 2;
            st.exchange_bytes += 8ull * (unsigned long long)ln * (unsigned long long)(P - 1);
            local_q = comm.h_all.ptr[me].out_len;

            PrRowSweepArgs<SizeT> rows;
            const bool by_rows = row_sweep && local_q > 0;
            if (by_rows) GR_TRY(BinRows(local_q, delta, threshold, src, &rows));
            st.by_rows = by_rows;

            // ---- (b) power iteration (pr_enactor.cuh:346-487) ---------------------------------------------
            GR_TRY(cudaEventRecord(ev1, stream));
            swept = true;
            int cgen = 0;
            for (;;) {
                GR_TRY(cudaMemsetAsync(Counters(), 0, sizeof(oprtr::LevelCounters), stream));
                if (by_rows) {
                    rows.contrib_in  = contrib.ptr + (unsigned long long)cgen * half;
                    rows.contrib_out = contrib.ptr + (unsigned long long)(cgen ^ 1) * half +
                                       (unsigned long long)me * (unsigned long long)ln;
                    if (row_tickets) {
                        const unsigned long long units = rows.bins.n_pieces + (rows.bins.n_medium + kMediumUnit - 1) / kMediumUnit +
                                                         (rows.bins.n_small + kSmallUnit - 1) / kSmallUnit;
                        const unsigned long long want = (units + oprtr::rows::kWarps - 1) / oprtr::rows::kWarps;
                        const unsigned long long resident = (unsigned long long)dev.sm_count * (unsigned long long)ticket_ctas_per_sm;
                        GR_TRY(cudaMemsetAsync(sweep_ticket.ptr, 0, sizeof(unsigned long long), stream));
                        PrRowSweepTicketKernel<SizeT><<<(unsigned)(want < resident ? (want ? want : 1) : resident),
                                                        oprtr::rows::kThreads, 0, stream>>>(rows, sweep_ticket.ptr);
                    } else {
                        PrRowSweepKernel<SizeT><<<rows.bins.Blocks(), oprtr::rows::kThreads, 0, stream>>>(rows);
                    }
                    GR_TRY(cudaGetLastError());
                    if (rows.bins.n_big) {
                        PrBigRowFinishKernel<SizeT><<<util::FillGrid(rows.bins.n_big), 256, 0, stream>>>(rows);
                        GR_TRY(cudaGetLastError());
                        ++enactor_stats.launches;
                    }
                } else {
                    if (local_q) GR_TRY(Advance<GatherFunctor>(MakeDataSlice(cgen, 0), local_q, st.iterations));
                    MgPrFinishKernel<<<util::FillGrid(row_grid), 256, 0, stream>>>(
                        ln, g.part, peers, cgen ^ 1, deg.ptr, rank.ptr, acc.ptr, delta, threshold, src, Counters());
                    GR_TRY(cudaGetLastError());
                }
                GR_TRY(comm.Exchange(Counters(), stream));
                enactor_stats.launches += 2;
                st.exchange_bytes += 8ull * (unsigned long long)ln * (unsigned long long)(P - 1);
                cgen ^= 1;
                ++st.iterations;
                unsigned long long moving = 0;
                for (int p = 0; p < P; ++p) { moving += comm.h_all.ptr[p].aux0; st.edges += comm.h_all.ptr[p].edges; }
                if (moving == 0 || st.iterations >= max_iteration) break;
            }
        }
        GR_TRY(cudaEventRecord(ev2, stream));
        GR_TRY(cudaStreamSynchronize(stream));
        float ms = 0;
        cudaEventElapsedTime(&ms, ev0, ev2);
        st.elapsed_ms = ms;
        if (swept) { cudaEventElapsedTime(&ms, ev1, ev2); st.sweep_ms = ms; }
        st.launches = enactor_stats.launches;
        if (out) *out = st;
        return cudaSuccess;
    }

    // fp64 ranks of the owned rows, local-row order
    cudaError_t Extract(double *h_rank)
    {
        GR_TRY(rank.ToHost(h_rank, (size_t)graph->local_nodes, stream));
        return cudaStreamSynchronize(stream);
    }
};

}  // namespace pr
}  // namespace app
}  // namespace gunrock
