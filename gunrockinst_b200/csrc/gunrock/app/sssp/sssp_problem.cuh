// sssp_problem.cuh -- per-run device state of single-source shortest paths.
//
// Counterpart of the reference's SSSPProblem (gunrock/app/sssp/sssp_problem.cuh:35-420):
//   distances  uint32, UINT_MAX = unreachable                         (:325)
//   preds      int32, initialised to the vertex itself                (:372)
//   weights    uint32 per CSR entry (the graph's edge_values)
//   delta      bucket width of the near/far queue: the reference intends
//              avg_edge_value * 32 / avg_degree * delta_factor (:377-381) but divides
//              0 by 0 on every path of this fork; here the two
//              averages are computed on the device at Init.
// MARK_PREDECESSORS is always true for the operators (they must hand the real source
// vertex to the functor, sssp_problem.cuh:42); MARK_PATHS selects whether parents are kept.
//
// With MARK_PATHS the distance and the parent of a vertex live in ONE 64-bit word
// (distance in the high half, parent in the low half) updated with a single 64-bit
// atomicMin.  A parent therefore always belongs to the distance next to it -- the
// reference writes the parent with a plain store after a 32-bit atomicMin
// (sssp_functor.cuh:52-84), which can pair a distance with the parent of a worse path --
// and the final parent is deterministic: the smallest vertex id among the tight in-edges.
#pragma once
#include "../problem_base.cuh"
#include "../../priority_queue/kernel.cuh"

namespace gunrock {
namespace app {
namespace sssp {

static __global__ void SumWeightsKernel(const unsigned *w, unsigned long long m, unsigned long long *out)
{
    __shared__ unsigned long long s_red[256 / 32 + 1];
    unsigned long long mine = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride)
        mine += w[i];
    unsigned long long total;
    (void)util::BlockExclusiveSum<256>(mine, s_red, total);
    if (threadIdx.x == 0 && total) atomicAdd(out, total);
}

// dist/pred words -> separate arrays (Extract)
static __global__ void UnpackKernel(const unsigned long long *packed, unsigned long long n, unsigned *dist, int *preds)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const unsigned long long p = packed[i];
        dist[i] = (unsigned)(p >> 32);
        preds[i] = (int)(unsigned)p;
    }
}

static __global__ void InitPackedKernel(unsigned long long *packed, unsigned long long n)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        packed[i] = (0xffffffffull << 32) | i;
}

template <typename _VertexId, typename _SizeT, bool _MARK_PATHS>
struct SSSPProblem : ProblemBase<_VertexId, _SizeT> {
    typedef _VertexId VertexId;
    typedef _SizeT    SizeT;
    typedef unsigned  Value;
    typedef ProblemBase<VertexId, SizeT> Base;
    static const bool MARK_PREDECESSORS  = true;      // operators pass the source vertex as s_id
    static const bool ENABLE_IDEMPOTENCE = false;
    static const bool MARK_PATHS         = _MARK_PATHS;

    struct DataSlice {
        unsigned           *d_labels;       // !MARK_PATHS: distances
        unsigned long long *d_dist_pred;    //  MARK_PATHS: (distance << 32) | parent
        const unsigned     *d_weights;
        unsigned           *d_mark;         // 1 bit / vertex: already in the frontier being built
        unsigned            delta;
        long long           nodes, local_nodes;
        graph::Partition    part;

        __device__ __forceinline__ unsigned Dist(VertexId v) const
        {
            if (MARK_PATHS) return (unsigned)(util::LoadCG(d_dist_pred + v) >> 32);
            return util::LoadCG(d_labels + v);
        }
        __host__ __device__ __forceinline__ unsigned long long BitPos(VertexId v) const { return (unsigned long long)v; }
    };

    util::DeviceArray<unsigned>           labels;
    util::DeviceArray<unsigned long long> dist_pred;
    util::DeviceArray<unsigned>           mark;
    util::DeviceArray<unsigned long long> scratch64;
    priority_queue::NearFarPile<VertexId> pile;
    unsigned delta = 1;
    double   avg_weight = 0, avg_degree = 0;
    VertexId src = 0;
    unsigned long long src_degree = 0;

    // reference: SSSPProblem::Init (sssp_problem.cuh:199-290)
    cudaError_t Init(typename Base::Graph *g, int delta_factor, cudaStream_t s)
    {
        GR_TRY(Base::InitBase(g, s));
        if (g->edge_values.size == 0 && g->edges > 0)
            return util::GRError(cudaErrorInvalidValue, "SSSP needs edge values", __FILE__, __LINE__);
        const size_t n = (size_t)g->local_nodes;
        if (MARK_PATHS) GR_TRY(dist_pred.Allocate(n)); else GR_TRY(labels.Allocate(n));
        GR_TRY(mark.Allocate((n + 31) / 32 + 1));
        GR_TRY(pile.Allocate(n));
        GR_TRY(scratch64.Allocate(1));
        GR_TRY(scratch64.Zero(s));
        unsigned long long sum = 0;
        if (g->edges) {
            SumWeightsKernel<<<util::FillGrid((unsigned long long)g->edges), 256, 0, s>>>(
                g->edge_values.ptr, (unsigned long long)g->edges, scratch64.ptr);
            GR_TRY(cudaGetLastError());
            GR_TRY(cudaMemcpyAsync(&sum, scratch64.ptr, sizeof(sum), cudaMemcpyDeviceToHost, s));
            GR_TRY(cudaStreamSynchronize(s));
        }
        avg_weight = g->edges ? (double)sum / (double)g->edges : 1.0;
        avg_degree = g->nodes ? (double)g->edges / (double)g->nodes : 1.0;
        if (delta_factor < 1) delta_factor = 1;
        double d = avg_degree > 0 ? avg_weight * 32.0 / avg_degree * (double)delta_factor : avg_weight;
        if (d < 1.0) d = 1.0;
        if (d > 4.0e9) d = 4.0e9;
        delta = (unsigned)d;
        return cudaSuccess;
    }

    DataSlice MakeDataSlice() const
    {
        DataSlice d;
        d.d_labels = labels.ptr; d.d_dist_pred = dist_pred.ptr;
        d.d_weights = this->graph->edge_values.ptr; d.d_mark = mark.ptr; d.delta = delta;
        d.nodes = this->graph->nodes; d.local_nodes = this->graph->local_nodes; d.part = this->graph->part;
        return d;
    }

    // reference: SSSPProblem::Reset (sssp_problem.cuh:300-400)
    cudaError_t Reset(VertexId source, double queue_sizing)
    {
        cudaStream_t s = this->stream;
        GR_TRY(Base::ResetBase(VERTEX_FRONTIERS, queue_sizing));
        src = source; src_degree = 0;
        const unsigned long long n = (unsigned long long)this->graph->local_nodes;
        if (MARK_PATHS) {
            if (n) { InitPackedKernel<<<util::FillGrid(n), 256, 0, s>>>(dist_pred.ptr, n); GR_TRY(cudaGetLastError()); }
        } else {
            if (n) GR_TRY(cudaMemsetAsync(labels.ptr, 0xff, sizeof(unsigned) * (size_t)n, s));
        }
        GR_TRY(mark.Zero(s));
        GR_TRY(pile.Reset(s));
        if (source < 0 || (long long)source >= this->graph->nodes) return cudaSuccess;
        if (MARK_PATHS) {
            const unsigned long long word = (unsigned long long)(unsigned)source;       // distance 0, parent = itself
            GR_TRY(cudaMemcpyAsync(dist_pred.ptr + source, &word, sizeof(word), cudaMemcpyHostToDevice, s));
            GR_TRY(cudaStreamSynchronize(s));
        } else {
            const unsigned zero = 0;
            GR_TRY(cudaMemcpyAsync(labels.ptr + source, &zero, sizeof(zero), cudaMemcpyHostToDevice, s));
            GR_TRY(cudaStreamSynchronize(s));
        }
        GR_TRY(cudaMemcpyAsync(this->frontier_queues[0].ptr, &source, sizeof(VertexId), cudaMemcpyHostToDevice, s));
        SizeT ends[2];
        GR_TRY(cudaMemcpyAsync(ends, this->graph->row_offsets.ptr + source, 2 * sizeof(SizeT), cudaMemcpyDeviceToHost, s));
        GR_TRY(cudaStreamSynchronize(s));
        src_degree = (unsigned long long)(ends[1] - ends[0]);
        return cudaSuccess;
    }

    // reference: SSSPProblem::Extract (sssp_problem.cuh:141-180)
    cudaError_t Extract(unsigned *h_dist, VertexId *h_preds)
    {
        cudaStream_t s = this->stream;
        const size_t n = (size_t)this->graph->local_nodes;
        if (!n) return cudaSuccess;
        if (MARK_PATHS) {
            // unpack into the two frontier queues (idle after Enact), then copy out
            unsigned *d_dist = reinterpret_cast<unsigned *>(this->frontier_queues[0].ptr);
            int *d_pred = reinterpret_cast<int *>(this->frontier_queues[1].ptr);
            UnpackKernel<<<util::FillGrid(n), 256, 0, s>>>(dist_pred.ptr, n, d_dist, d_pred);
            GR_TRY(cudaGetLastError());
            if (h_dist)  GR_TRY(cudaMemcpyAsync(h_dist, d_dist, sizeof(unsigned) * n, cudaMemcpyDeviceToHost, s));
            if (h_preds) GR_TRY(cudaMemcpyAsync(h_preds, d_pred, sizeof(int) * n, cudaMemcpyDeviceToHost, s));
        } else {
            if (h_dist) GR_TRY(labels.ToHost(h_dist, n, s));
        }
        return cudaStreamSynchronize(s);
    }
};

}  // namespace sssp
}  // namespace app
}  // namespace gunrock
