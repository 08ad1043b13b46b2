// mgcc_enactor.cuh -- 1D vertex-partitioned connected components, one device per process,
// label updates delivered through peer memory (NVLink / NVSwitch).
//
// No reference counterpart for the partitioned part (gunrock/app/problem_base.cuh:336-338 is a
// TODO).  The output convention is the single-device CC's (gunrock/app/cc/cc_problem.cuh:164-170,
// 262-272 as restated in cc_enactor.cuh of this directory): every vertex ends with the smallest
// vertex id of its component, an isolated vertex with its own id.
//
// The single-device enactor hooks and pointer-jumps over the edge list; both passes chase labels
// of arbitrary vertices, i.e. remote reads on a partitioned graph.  The partitioned enactor
// propagates minima instead, on the machinery of the partitioned SSSP (sssp/mgsssp_enactor.cuh):
// every owned row has one 64-bit word label << 32 | vertex that ANY device may lower with a
// 64-bit atomicMin (a fire-and-forget reduction over NVLink when the row lives on a peer), a
// 4-byte-per-global-vertex replica of the best label already sent culls updates that cannot
// improve anything, and the owner collects touched rows whose label dropped into its next
// frontier.  All rows start in the frontier with their own id; the loop ends when no device holds
// an unsent improvement, after O(diameter) iterations.  Labels are exact whatever the
// interleaving: every stored value is the id of a vertex of the same component, and the minimum
// of the component reaches every member before the frontier can empty.  Needs a symmetric graph
// (labels travel along out-edges only), like the partitioned BFS.
#pragma once
#include "../sssp/mgsssp_enactor.cuh"

namespace gunrock {
namespace app {
namespace cc {

template <typename VertexId, typename SizeT, typename ProblemData>
struct MgCcFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *p, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        const unsigned label = (unsigned)(util::LoadCG(p->own + p->part.Local(s_id)) >> 32);
        const int owner = p->part.Owner(d_id);
        const long long row = p->part.Local(d_id);
        unsigned *b = p->best + ((unsigned long long)owner * (unsigned long long)p->local_nodes + (unsigned long long)row);
        if (label >= util::LoadCG(b)) return false;
        if (label >= atomicMin(b, label)) return false;
        const unsigned long long word = ((unsigned long long)label << 32) | (unsigned long long)(unsigned)d_id;
        atomicMin(p->peers.dist_pred[owner] + row, word);                       // result unused: a reduction
        atomicOr(p->peers.touched[owner] + (row >> 5), 1u << (row & 31));
        return false;                                                           // the owner builds the frontier
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *p, SizeT e_id = 0, SizeT e_id_in = 0) {}
    static __device__ __forceinline__ bool CondFilter(VertexId node, const DataSlice *p, unsigned v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(VertexId node, const DataSlice *p, unsigned v = 0, SizeT nid = 0) {}
};

// every owned row: label = own id, queued with it, member of the first frontier
static __global__ void MgCcInitKernel(unsigned long long *label_word, unsigned *queued, int *queue, long long local_nodes,
                                      graph::Partition part)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < local_nodes; r += stride) {
        const unsigned v = (unsigned)part.Global(r);
        label_word[r] = ((unsigned long long)v << 32) | (unsigned long long)v;
        queued[r] = v;
        queue[r] = (int)v;
    }
}

// replica of the best label sent per GLOBAL vertex (owner-major): its own id
static __global__ void MgCcInitBestKernel(unsigned *best, long long local_nodes, graph::Partition part)
{
    const long long total = local_nodes << part.log_parts;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long pos = (long long)blockIdx.x * blockDim.x + threadIdx.x; pos < total; pos += stride) {
        graph::Partition of = part;
        of.rank = (int)(pos / local_nodes);
        best[pos] = (unsigned)of.Global(pos % local_nodes);
    }
}

template <typename _SizeT>
class MGCCEnactor : public EnactorBase {
  public:
    typedef _SizeT SizeT;
    typedef int    VertexId;
    typedef sssp::MgSsspProblem<SizeT> Problem;          // same device state: 64-bit words, best replica, peer table
    typedef typename Problem::DataSlice DataSlice;
    typedef MgCcFunctor<VertexId, SizeT, Problem> Functor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, false> AdvancePolicy;
    typedef graph::DeviceCsr<VertexId, SizeT> Graph;

    struct RunStats {
        double elapsed_ms = 0;
        long long iterations = 0, launches = 0;
        unsigned long long updates = 0, queued = 0;
    };

    Graph *graph = nullptr;
    util::DeviceArray<unsigned long long> label_word;
    util::DeviceArray<unsigned> queued, best, touched;
    util::DeviceArray<VertexId> queue;
    oprtr::advance::LbScratch<SizeT> lb_scratch;
    bfs::PeerComm comm;
    sssp::SsspPeerTable peers;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;

    ~MGCCEnactor()
    {
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
    }
    oprtr::LevelCounters *Counters() { return work_progress.Slot(1); }

    cudaError_t Setup(Graph *g, cudaStream_t s)
    {
        graph = g;
        if (!g->symmetric)
            return util::GRError(cudaErrorInvalidValue, "partitioned CC needs a symmetric graph", __FILE__, __LINE__);
        if (g->local_nodes % 32 != 0 && g->part.log_parts != 0)
            return util::GRError(cudaErrorInvalidValue, "partition segments must be word aligned", __FILE__, __LINE__);
        GR_TRY(SetupBase(g->device, s));
        const size_t ln = (size_t)g->local_nodes;
        GR_TRY(label_word.Allocate(ln + 1));
        GR_TRY(queued.Allocate(ln + 1));
        GR_TRY(best.Allocate(ln * (size_t)g->part.Parts() + 1));
        GR_TRY(touched.Allocate((ln + 31) / 32 + 1));
        GR_TRY(queue.Allocate(ln + 1));
        GR_TRY(lb_scratch.Allocate(ln + 1, (size_t)g->edges + 1, AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        GR_TRY(comm.Setup(g->part.Parts(), g->part.rank, s));
        peers = sssp::SsspPeerTable();
        peers.dist_pred[g->part.rank] = label_word.ptr;
        peers.touched[g->part.rank] = touched.ptr;
        GR_TRY(cudaEventCreate(&ev0));
        GR_TRY(cudaEventCreate(&ev1));
        return cudaStreamSynchronize(s);
    }

    DataSlice MakeDataSlice() const
    {
        DataSlice d;
        d.own = label_word.ptr; d.best = best.ptr; d.d_weights = nullptr;
        d.nodes = graph->nodes; d.local_nodes = graph->local_nodes; d.part = graph->part;
        d.peers = peers;
        return d;
    }

    // every rank calls it
    cudaError_t Run(RunStats *out)
    {
        if (!comm.ready) return util::GRError(cudaErrorNotReady, "peer mappings have not been imported", __FILE__, __LINE__);
        Graph &g = *graph;
        const long long ln = g.local_nodes;
        GR_TRY(cudaEventRecord(ev0, stream));
        GR_TRY(ResetBase());
        MgCcInitKernel<<<util::FillGrid((unsigned long long)ln + 1), 256, 0, stream>>>(label_word.ptr, queued.ptr, queue.ptr, ln, g.part);
        GR_TRY(cudaGetLastError());
        MgCcInitBestKernel<<<util::FillGrid((unsigned long long)(ln << g.part.log_parts) + 1), 256, 0, stream>>>(best.ptr, ln, g.part);
        GR_TRY(cudaGetLastError());
        GR_TRY(touched.Zero(stream));
        oprtr::LevelCounters first = {};
        first.out_len = (unsigned long long)ln;
        first.out_degree = (unsigned long long)g.edges;
        GR_TRY(cudaMemcpyAsync(Counters(), &first, sizeof(first), cudaMemcpyHostToDevice, stream));
        GR_TRY(cudaStreamSynchronize(stream));          // `first` lives on this frame
        GR_TRY(comm.Exchange(Counters(), stream));     // also the barrier behind every device's reset
        enactor_stats.launches += 3;

        const int P = comm.table.parts, me = comm.table.rank;
        unsigned long long frontier = 0, updates = 0, total_queued = 0;
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
            GR_TRY(comm.Barrier(stream));               // every remote update of this iteration has landed
            sssp::MgSsspCollectKernel<SizeT, 256><<<util::FillGrid((unsigned long long)((ln + 31) >> 5)), 256, 0, stream>>>(
                touched.ptr, label_word.ptr, queued.ptr, queue.ptr, (unsigned long long)queue.size,
                g.row_offsets.ptr, ln, g.part, Counters());
            GR_TRY(cudaGetLastError());
            GR_TRY(comm.Exchange(Counters(), stream));
            enactor_stats.launches += 3;
            total_queued += frontier;
            frontier = 0;
            unsigned long long ovf = 0;
            for (int p = 0; p < P; ++p) {
                frontier += comm.h_all.ptr[p].out_len;
                updates += comm.h_all.ptr[p].edges;
                ovf += comm.h_all.ptr[p].overflow;
            }
            if (ovf) return util::GRError(cudaErrorInvalidConfiguration, "frontier queue overflow on a partitioned CC iteration", __FILE__, __LINE__);
            local_q = comm.h_all.ptr[me].out_len; local_hint = comm.h_all.ptr[me].out_degree;
            ++iteration;
        }
        GR_TRY(cudaEventRecord(ev1, stream));
        GR_TRY(cudaEventSynchronize(ev1));
        float ms = 0;
        cudaEventElapsedTime(&ms, ev0, ev1);
        if (out) {
            out->elapsed_ms = ms; out->iterations = iteration; out->launches = enactor_stats.launches;
            out->updates = updates; out->queued = total_queued;
        }
        return cudaSuccess;
    }

    // labels of the owned rows in local-row order; *roots = owned vertices that are their component's label
    cudaError_t Extract(int *h_labels, long long *roots)
    {
        const size_t ln = (size_t)graph->local_nodes;
        std::vector<unsigned long long> h(ln ? ln : 1);
        GR_TRY(label_word.ToHost(h.data(), ln, stream));
        GR_TRY(cudaStreamSynchronize(stream));
        long long r = 0;
        for (size_t i = 0; i < ln; ++i) {
            const int label = (int)(unsigned)(h[i] >> 32);
            if (h_labels) h_labels[i] = label;
            r += (unsigned)label == (unsigned)graph->part.Global((long long)i);
        }
        if (roots) *roots = r;
        return cudaSuccess;
    }
};

}  // namespace cc
}  // namespace app
}  // namespace gunrock
