// pr_problem.cuh -- device state of PageRank.
//
// Counterpart of the reference's PRProblem (gunrock/app/pr/pr_problem.cuh:35-470):
//   d_rank_curr / d_rank_next   ranks of this and the next iteration; curr starts at 1 - delta
//                               on every vertex (:407), next at 0 (the reference relies on fresh
//                               cudaMalloc memory for that, :337-343; here it is zeroed)
//   d_degrees / d_degrees_pong  out-degree, iteratively reduced while zero-out-degree vertices
//                               are pruned; a pruned vertex ends at -1 (pr_functor.cuh:113-175)
//   d_node_ids                  iota, permuted by the final sort (:413)
//   delta, threshold, src_node  by value in the DataSlice (the reference keeps 1-element device
//                               arrays, :86-88)
// MARK_PREDECESSORS is true so that the operators hand the functor the real source (:42).
//
// Ranks are held in fp64 (the reference: fp32 + float atomicAdd, order dependent): the
// un-normalised rank of an R-MAT hub reaches 1e5, where fp32 resolves only ~1e-2 -- the
// per-vertex 1e-6 agreement with the fp64 oracle needs the wider type.  The fp32 view the C
// API returns is a final conversion.  d_contrib caches rank_curr / degree so that an edge costs
// one 8-byte read instead of two reads and an fp64 division.
#pragma once
#include "../problem_base.cuh"
#include "../../util/radix_sort.cuh"

namespace gunrock {
namespace app {
namespace pr {

template <typename SizeT>
__global__ void InitDegreesKernel(const SizeT *row_offsets, unsigned long long n, int *deg, int *pong,
                                  double *curr, double *next, double *contrib, double init)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long v = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; v < n; v += stride) {
        const int d = (int)(row_offsets[v + 1] - row_offsets[v]);
        deg[v] = d; pong[v] = d;
        curr[v] = init; next[v] = 0.0;
        contrib[v] = 0.0;
    }
}

// contrib = curr / degree for the vertices that survived the pruning
static __global__ void InitContribKernel(const int *deg, const double *curr, double *contrib, unsigned long long n)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long v = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; v < n; v += stride)
        contrib[v] = deg[v] > 0 ? curr[v] / (double)deg[v] : 0.0;
}

// sort key of vertex v: descending rank, ties by ascending id (a stable descending sort of
// (rank, id) pairs, which is what the reference's CUB call produces, pr_enactor.cuh:497-500)
static __global__ void MakeSortKeysKernel(const double *rank, unsigned long long n, unsigned long long *keys)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long v = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; v < n; v += stride) {
        unsigned bits = __float_as_uint((float)rank[v]);
        bits = (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);    // ascending order-preserving
        keys[v] = ((unsigned long long)(~bits) << 32) | v;             // ~ => descending
    }
}

static __global__ void UnpackSortedKernel(const unsigned long long *keys, unsigned long long n, int *ids, float *ranks)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const unsigned long long k = keys[i];
        unsigned bits = ~(unsigned)(k >> 32);
        bits = (bits & 0x80000000u) ? (bits & 0x7fffffffu) : ~bits;
        ids[i] = (int)(unsigned)k;
        ranks[i] = __uint_as_float(bits);
    }
}

template <typename _VertexId, typename _SizeT, typename _Value>
struct PRProblem : ProblemBase<_VertexId, _SizeT> {
    typedef _VertexId VertexId;
    typedef _SizeT    SizeT;
    typedef _Value    Value;                          // double
    typedef ProblemBase<VertexId, SizeT> Base;
    static const bool MARK_PREDECESSORS  = true;
    static const bool ENABLE_IDEMPOTENCE = false;

    struct DataSlice {
        Value    *d_rank_curr, *d_rank_next, *d_contrib;
        int      *d_degrees, *d_degrees_pong;
        Value     delta, threshold;
        VertexId  src_node;
        long long nodes, local_nodes;
        graph::Partition part;
    };

    util::DeviceArray<Value> rank[2], contrib;
    util::DeviceArray<int>   degrees, degrees_pong;
    util::DeviceArray<unsigned long long> sort_a, sort_b;
    util::RadixSortScratch   sort_scratch;
    int      curr = 0;                                // rank[curr] is d_rank_curr
    Value    delta = 0.85, threshold = 0.01;
    VertexId src_node = -1;

    cudaError_t Init(typename Base::Graph *g, cudaStream_t s)
    {
        GR_TRY(Base::InitBase(g, s));
        const size_t n = (size_t)g->local_nodes;
        GR_TRY(rank[0].Allocate(n)); GR_TRY(rank[1].Allocate(n)); GR_TRY(contrib.Allocate(n));
        GR_TRY(degrees.Allocate(n)); GR_TRY(degrees_pong.Allocate(n));
        return cudaSuccess;
    }
    DataSlice MakeDataSlice() const
    {
        DataSlice d;
        d.d_rank_curr = rank[curr].ptr; d.d_rank_next = rank[curr ^ 1].ptr; d.d_contrib = contrib.ptr;
        d.d_degrees = degrees.ptr; d.d_degrees_pong = degrees_pong.ptr;
        d.delta = delta; d.threshold = threshold; d.src_node = src_node;
        d.nodes = this->graph->nodes; d.local_nodes = this->graph->local_nodes; d.part = this->graph->part;
        return d;
    }
    // reference: PRProblem::Reset (pr_problem.cuh:300-440)
    cudaError_t Reset(VertexId src, Value delta_, Value threshold_)
    {
        GR_TRY(Base::ResetBase(VERTEX_FRONTIERS, 1.0));       // queue_sizing fixed at 1.0 (:318)
        src_node = src; delta = delta_; threshold = threshold_; curr = 0;
        const unsigned long long n = (unsigned long long)this->graph->local_nodes;
        if (!n) return cudaSuccess;
        InitDegreesKernel<SizeT><<<util::FillGrid(n), 256, 0, this->stream>>>(
            this->graph->row_offsets.ptr, n, degrees.ptr, degrees_pong.ptr, rank[0].ptr, rank[1].ptr,
            contrib.ptr, (Value)(1.0 - delta));
        return util::GRError(cudaGetLastError(), "InitDegreesKernel failed", __FILE__, __LINE__);
    }
    // ranks in vertex order (fp64) and / or the reference's output: all (id, fp32 rank) pairs in
    // descending rank (pr_problem.cuh:150-176 after the sort of pr_enactor.cuh:497-500)
    cudaError_t Extract(Value *h_rank64, VertexId *h_sorted_ids, float *h_sorted_ranks, size_t top = 0)
    {
        cudaStream_t s = this->stream;
        const size_t n = (size_t)this->graph->local_nodes;
        if (!n) return cudaSuccess;
        if (h_rank64) GR_TRY(rank[curr].ToHost(h_rank64, n, s));
        if (h_sorted_ids || h_sorted_ranks) {
            GR_TRY(sort_a.Allocate(n)); GR_TRY(sort_b.Allocate(n));
            MakeSortKeysKernel<<<util::FillGrid(n), 256, 0, s>>>(rank[curr].ptr, n, sort_a.ptr);
            GR_TRY(cudaGetLastError());
            unsigned long long *sorted = nullptr;
            GR_TRY(util::RadixSort64(sort_a.ptr, sort_b.ptr, n, 64, sort_scratch, &sorted, s));
            // unpack into the (idle) frontier queue and the pong degrees
            int *d_ids = reinterpret_cast<int *>(this->frontier_queues[0].ptr);
            float *d_r = reinterpret_cast<float *>(degrees_pong.ptr);
            UnpackSortedKernel<<<util::FillGrid(n), 256, 0, s>>>(sorted, n, d_ids, d_r);
            GR_TRY(cudaGetLastError());
            const size_t k = (top && top < n) ? top : n;
            if (h_sorted_ids)   GR_TRY(cudaMemcpyAsync(h_sorted_ids, d_ids, sizeof(int) * k, cudaMemcpyDeviceToHost, s));
            if (h_sorted_ranks) GR_TRY(cudaMemcpyAsync(h_sorted_ranks, d_r, sizeof(float) * k, cudaMemcpyDeviceToHost, s));
        }
        return cudaStreamSynchronize(s);
    }
};

}  // namespace pr
}  // namespace app
}  // namespace gunrock
