// mgsssp_enactor.cuh -- 1D vertex-partitioned single-source shortest paths, one device per
// process, remote relaxations delivered through peer memory (NVLink / NVSwitch).
//
// No reference counterpart for the partitioned part (gunrock/app/problem_base.cuh:336-338 is a
// TODO); the per-edge arithmetic is the reference's relaxation (gunrock/app/sssp/
// sssp_functor.cuh:52-66: new = dist[s] + w, unsigned, atomicMin on the destination) and the
// outputs follow SSSPProblem (sssp_problem.cuh:325,372): uint32 distances, UINT_MAX =
// unreachable, predecessor = own id for the source and for unreached vertices.
//
// Every device holds, for the rows it owns, one 64-bit word (distance << 32 | predecessor) that
// ANY device may lower with a 64-bit atomicMin -- a remote one travels as a fire-and-forget
// reduction over NVLink, straight from the advance kernel that found the shorter path, together
// with an atomicOr on the owner's "touched" bitmap.  A full-size replica of the best distance
// this device has ever sent per vertex (`best`, 4 bytes per GLOBAL vertex, owner-major index)
// culls relaxations that cannot improve anything before they reach the link.  An iteration is
//     advance   relax the out-edges of the local frontier (LB / TWC operator, no output queue)
//     barrier   flag words in the peers' mailboxes (bfs/mgbfs_p2p.cuh)
//     collect   owner scans its touched words: a vertex whose distance is now below the value it
//               was last queued with joins the next local frontier
//     exchange  64-byte counter blocks of all devices -> global frontier length (termination)
// i.e. a frontier-driven Bellman-Ford; distances are exact whatever the interleaving because
// every stored value is the length of a real path and the loop ends only when no device holds
// an un-relaxed improvement.
#pragma once
#include <vector>
#include "../enactor_base.cuh"
#include "../bfs/mgbfs_p2p.cuh"
#include "../enactor_base.cuh"
#include "../../util/memset_kernel.cuh"

namespace gunrock {
namespace app {
namespace sssp {

struct SsspPeerTable {
    unsigned long long *dist_pred[bfs::kMaxPeers] = {};     // [local_nodes] of every device
    unsigned           *touched[bfs::kMaxPeers]   = {};     // [local_nodes / 32] of every device
};

template <typename _SizeT>
struct MgSsspProblem {
    typedef int      VertexId;
    typedef _SizeT   SizeT;
    typedef unsigned Value;
    static const bool MARK_PREDECESSORS  = true;    // the operators hand the real source vertex to the functor
    static const bool ENABLE_IDEMPOTENCE = false;
    struct DataSlice {
        const unsigned long long *own;      // this device's dist|pred words
        unsigned       *best;               // replica: best distance sent per global vertex
        const unsigned *d_weights;
        long long       nodes, local_nodes;
        graph::Partition part;
        SsspPeerTable   peers;
    };
};

template <typename VertexId, typename SizeT, typename ProblemData>
struct MgSsspFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *p, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        const unsigned label = (unsigned)(util::LoadCG(p->own + p->part.Local(s_id)) >> 32);
        const unsigned nd = label + util::LoadStream(p->d_weights + e_id);      // unsigned, no guard (reference)
        const int owner = p->part.Owner(d_id);
        const long long row = p->part.Local(d_id);
        unsigned *b = p->best + ((unsigned long long)owner * (unsigned long long)p->local_nodes + (unsigned long long)row);
        if (nd >= util::LoadCG(b)) return false;
        if (nd >= atomicMin(b, nd)) return false;
        const unsigned long long word = ((unsigned long long)nd << 32) | (unsigned long long)(unsigned)s_id;
        atomicMin(p->peers.dist_pred[owner] + row, word);                       // result unused: a reduction
        atomicOr(p->peers.touched[owner] + (row >> 5), 1u << (row & 31));
        return false;                                                           // the owner builds the frontier
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *p, SizeT e_id = 0, SizeT e_id_in = 0) {}
    static __device__ __forceinline__ bool CondFilter(VertexId node, const DataSlice *p, unsigned v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(VertexId node, const DataSlice *p, unsigned v = 0, SizeT nid = 0) {}
};

// dist|pred <- UINT_MAX|own id (sssp_problem.cuh:325,372), queued <- UINT_MAX
static __global__ void MgSsspInitKernel(unsigned long long *dist_pred, unsigned *queued, long long local_nodes,
                                        graph::Partition part)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < local_nodes; r += stride) {
        dist_pred[r] = 0xffffffff00000000ull | (unsigned long long)(unsigned)part.Global(r);
        queued[r] = 0xffffffffu;
    }
}

template <typename SizeT>
__global__ void MgSsspSeedKernel(unsigned long long *dist_pred, unsigned *queued, unsigned *best, int *queue,
                                 oprtr::LevelCounters *c, const SizeT *row_offsets, long long local_nodes,
                                 graph::Partition part, int src, long long nodes)
{
    c->out_len = 0; c->out_degree = 0; c->edges = 0; c->scan_total = 0; c->aux0 = 0; c->aux1 = 0; c->overflow = 0; c->reserved = 0;
    if (src < 0 || (long long)src >= nodes) return;
    const long long row = part.Local(src);
    best[(unsigned long long)part.Owner(src) * (unsigned long long)local_nodes + (unsigned long long)row] = 0;
    if (part.Mine(src)) {
        dist_pred[row] = (unsigned long long)(unsigned)src;         // distance 0, predecessor = itself
        queued[row] = 0;
        queue[0] = src;
        c->out_len = 1;
        c->out_degree = (unsigned long long)(row_offsets[row + 1] - row_offsets[row]);
    }
}

// owner side of an iteration: touched vertices whose distance dropped below the value they were
// last queued with form the next local frontier (CTA-aggregated enqueue)
template <typename SizeT, int THREADS>
__global__ void __launch_bounds__(THREADS)
MgSsspCollectKernel(unsigned *touched, const unsigned long long *dist_pred, unsigned *queued, int *queue,
                    unsigned long long cap, const SizeT *row_offsets, long long local_nodes,
                    graph::Partition part, oprtr::LevelCounters *slot)
{
    __shared__ unsigned s_warp[THREADS / 32 + 1];
    __shared__ unsigned long long s_red[THREADS / 32 + 1];
    __shared__ unsigned long long s_base;
    const long long words = (local_nodes + 31) >> 5;
    const long long rounds = (words + (long long)gridDim.x * THREADS - 1) / ((long long)gridDim.x * THREADS);
    unsigned long long my_degree = 0;
    for (long long r = 0; r < rounds; ++r) {
        const long long w = (r * gridDim.x + blockIdx.x) * THREADS + threadIdx.x;
        unsigned bits = 0;
        if (w < words) {
            unsigned t = util::LoadCG(touched + w);
            if (t) touched[w] = 0;
            while (t) {
                const int b = __ffs(t) - 1;
                t &= t - 1;
                const long long row = (w << 5) + b;
                const unsigned d = (unsigned)(util::LoadCG(dist_pred + row) >> 32);
                if (d < queued[row]) { queued[row] = d; bits |= 1u << b; }
            }
        }
        unsigned total;
        const unsigned excl = util::BlockExclusiveSum<THREADS>((unsigned)__popc(bits), s_warp, total);
        if (total == 0) continue;                       // uniform
        if (threadIdx.x == 0) s_base = atomicAdd(&slot->out_len, (unsigned long long)total);
        __syncthreads();
        unsigned long long at = s_base + excl;
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            const long long row = (w << 5) + b;
            my_degree += (unsigned long long)(row_offsets[row + 1] - row_offsets[row]);
            if (at < cap) queue[at] = (int)part.Global(row); else slot->overflow = 1;
            ++at;
        }
        __syncthreads();
    }
    oprtr::advance::BlockAccumulate<THREADS>(my_degree, s_red, &slot->out_degree);
}

template <typename _SizeT>
class MGSSSPEnactor : public EnactorBase {
  public:
    typedef _SizeT SizeT;
    typedef int    VertexId;
    typedef MgSsspProblem<SizeT> Problem;
    typedef typename Problem::DataSlice DataSlice;
    typedef MgSsspFunctor<VertexId, SizeT, Problem> Functor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, false> AdvancePolicy;
    typedef graph::DeviceCsr<VertexId, SizeT> Graph;

    struct SearchStats {
        double elapsed_ms = 0;
        long long iterations = 0, launches = 0;
        unsigned long long relaxations = 0, queued = 0;
    };

    Graph *graph = nullptr;
    util::DeviceArray<unsigned long long> dist_pred;
    util::DeviceArray<unsigned> queued, best, touched;
    util::DeviceArray<VertexId> queue;
    oprtr::advance::LbScratch<SizeT> lb_scratch;
    bfs::PeerComm comm;
    SsspPeerTable peers;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;

    ~MGSSSPEnactor()
    {
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
    }
    oprtr::LevelCounters *Counters() { return work_progress.Slot(1); }

    cudaError_t Setup(Graph *g, cudaStream_t s)
    {
        graph = g;
        if (!g->edge_values.size && g->edges)
            return util::GRError(cudaErrorInvalidValue, "partitioned SSSP needs edge weights", __FILE__, __LINE__);
        if (g->local_nodes % 32 != 0 && g->part.log_parts != 0)
            return util::GRError(cudaErrorInvalidValue, "partition segments must be word aligned", __FILE__, __LINE__);
        GR_TRY(SetupBase(g->device, s));
        const size_t ln = (size_t)g->local_nodes;
        GR_TRY(dist_pred.Allocate(ln + 1));
        GR_TRY(queued.Allocate(ln + 1));
        GR_TRY(best.Allocate(ln * (size_t)g->part.Parts() + 1));
        GR_TRY(touched.Allocate((ln + 31) / 32 + 1));
        GR_TRY(queue.Allocate(ln + 1));
        GR_TRY(lb_scratch.Allocate(ln + 1, (size_t)g->edges + 1, AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        GR_TRY(comm.Setup(g->part.Parts(), g->part.rank, s));
        peers = SsspPeerTable();
        peers.dist_pred[g->part.rank] = dist_pred.ptr;
        peers.touched[g->part.rank] = touched.ptr;
        GR_TRY(cudaEventCreate(&ev0));
        GR_TRY(cudaEventCreate(&ev1));
        return cudaStreamSynchronize(s);
    }

    DataSlice MakeDataSlice() const
    {
        DataSlice d;
        d.own = dist_pred.ptr; d.best = best.ptr; d.d_weights = graph->edge_values.ptr;
        d.nodes = graph->nodes; d.local_nodes = graph->local_nodes; d.part = graph->part;
        d.peers = peers;
        return d;
    }

    // every rank calls it with the same source
    cudaError_t Run(VertexId src, SearchStats *out)
    {
        if (!comm.ready) return util::GRError(cudaErrorNotReady, "peer mappings have not been imported", __FILE__, __LINE__);
        Graph &g = *graph;
        const long long ln = g.local_nodes;
        GR_TRY(cudaEventRecord(ev0, stream));
        GR_TRY(ResetBase());
        MgSsspInitKernel<<<util::FillGrid((unsigned long long)ln + 1), 256, 0, stream>>>(dist_pred.ptr, queued.ptr, ln, g.part);
        GR_TRY(cudaGetLastError());
        GR_TRY(cudaMemsetAsync(best.ptr, 0xff, sizeof(unsigned) * best.size, stream));
        GR_TRY(touched.Zero(stream));
        MgSsspSeedKernel<SizeT><<<1, 1, 0, stream>>>(dist_pred.ptr, queued.ptr, best.ptr, queue.ptr, Counters(),
                                                     g.row_offsets.ptr, ln, g.part, src, g.nodes);
        GR_TRY(cudaGetLastError());
        enactor_stats.launches += 2;
        GR_TRY(comm.Exchange(Counters(), stream));     // also the barrier behind every device's reset
        ++enactor_stats.launches;

        const int P = comm.table.parts, me = comm.table.rank;
        unsigned long long frontier = 0, relaxations = 0, total_queued = 0;
        for (int p = 0; p < P; ++p) frontier += comm.h_all.ptr[p].out_len;
        unsigned long long local_q = comm.h_all.ptr[me].out_len, local_hint = comm.h_all.ptr[me].out_degree;
        long long iteration = 0;
        const DataSlice ds = MakeDataSlice();
        while (frontier > 0) {
            GR_TRY(cudaMemsetAsync(Counters(), 0, sizeof(oprtr::LevelCounters), stream));
            if (local_q > 0) {
                oprtr::advance::Args<VertexId, SizeT> a;
                a.g = g.View();
                a.in_queue = queue.ptr; a.out_queue = nullptr;
                a.in_len = local_q; a.d_in_len = nullptr; a.out_cap = 0;
                a.slots = work_progress.d_slots.ptr; a.queue_index = 0; a.iteration = iteration;
                a.output = oprtr::advance::OUT_NONE;
                const oprtr::advance::MODE mode = local_hint >= 4096 ? oprtr::advance::LB : oprtr::advance::TWC_FORWARD;
                enactor_stats.launches += (mode == oprtr::advance::LB) ? 2 : 1;
                GR_TRY((oprtr::advance::LaunchKernel<AdvancePolicy, Problem, Functor>(mode, dev, ds, a, &lb_scratch,
                                                                                     local_hint, stream)));
            }
            GR_TRY(comm.Barrier(stream));               // every remote relaxation of this iteration has landed
            MgSsspCollectKernel<SizeT, 256><<<util::FillGrid((unsigned long long)((ln + 31) >> 5)), 256, 0, stream>>>(
                touched.ptr, dist_pred.ptr, queued.ptr, queue.ptr, (unsigned long long)queue.size,
                g.row_offsets.ptr, ln, g.part, Counters());
            GR_TRY(cudaGetLastError());
            GR_TRY(comm.Exchange(Counters(), stream));
            enactor_stats.launches += 3;
            total_queued += frontier;
            frontier = 0;
            unsigned long long ovf = 0;
            for (int p = 0; p < P; ++p) {
                frontier += comm.h_all.ptr[p].out_len;
                relaxations += comm.h_all.ptr[p].edges;
                ovf += comm.h_all.ptr[p].overflow;
            }
            if (ovf) return util::GRError(cudaErrorInvalidConfiguration, "frontier queue overflow on a partitioned SSSP iteration", __FILE__, __LINE__);
            local_q = comm.h_all.ptr[me].out_len; local_hint = comm.h_all.ptr[me].out_degree;
            ++iteration;
        }
        GR_TRY(cudaEventRecord(ev1, stream));
        GR_TRY(cudaEventSynchronize(ev1));
        float ms = 0;
        cudaEventElapsedTime(&ms, ev0, ev1);
        if (out) {
            out->elapsed_ms = ms; out->iterations = iteration; out->launches = enactor_stats.launches;
            out->relaxations = relaxations; out->queued = total_queued;
        }
        return cudaSuccess;
    }

    // distances / predecessors of the owned rows, in local-row order
    cudaError_t Extract(unsigned *h_dist, int *h_preds)
    {
        const size_t ln = (size_t)graph->local_nodes;
        std::vector<unsigned long long> h(ln ? ln : 1);
        GR_TRY(dist_pred.ToHost(h.data(), ln, stream));
        GR_TRY(cudaStreamSynchronize(stream));
        for (size_t i = 0; i < ln; ++i) {
            if (h_dist) h_dist[i] = (unsigned)(h[i] >> 32);
            if (h_preds) h_preds[i] = (int)(unsigned)(h[i] & 0xffffffffull);
        }
        return cudaSuccess;
    }
};

}  // namespace sssp
}  // namespace app
}  // namespace gunrock
