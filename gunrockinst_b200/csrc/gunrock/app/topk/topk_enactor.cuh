// topk_enactor.cuh -- top-k vertices by degree centrality (in-degree + out-degree).
//
// Counterpart of the reference's TOPKProblem / TOPKEnactor (gunrock/app/topk/topk_problem.cuh,
// topk_enactor.cuh:118-205): there, three vector kernels form degree_sum = out + in and three
// CUB radix sorts (descending, all keyed by the same sum) carry along the vertex ids, the
// in-degrees and the out-degrees.  Here ONE stable LSD radix sort (util/radix_sort.cuh) of the
// packed key (~degree_sum << 32 | vertex id) yields the same order -- descending sum, ties in
// ascending vertex id, which is what a stable descending sort of (sum, iota) produces -- and the
// k winners' degrees are gathered afterwards.
#pragma once
#include "../../graph/csr.cuh"
#include "../../util/radix_sort.cuh"
#include "../../util/memset_kernel.cuh"

namespace gunrock {
namespace app {
namespace topk {

template <typename SizeT>
__global__ void DegreeKeysKernel(const SizeT *out_offsets, const SizeT *in_offsets, long long n, unsigned long long *keys)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < n; v += stride) {
        const unsigned out_deg = (unsigned)(out_offsets[v + 1] - out_offsets[v]);
        const unsigned in_deg  = (unsigned)(in_offsets[v + 1] - in_offsets[v]);
        keys[v] = ((unsigned long long)(~(out_deg + in_deg)) << 32) | (unsigned long long)(unsigned)v;
    }
}

template <typename SizeT>
__global__ void GatherTopKernel(const unsigned long long *sorted, const SizeT *out_offsets, const SizeT *in_offsets,
                                int k, int *node_ids, int *in_degrees, int *out_degrees)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k) return;
    const unsigned v = (unsigned)(sorted[i] & 0xffffffffull);
    node_ids[i] = (int)v;
    in_degrees[i]  = (int)(in_offsets[v + 1] - in_offsets[v]);
    out_degrees[i] = (int)(out_offsets[v + 1] - out_offsets[v]);
}

// out_offsets / in_offsets: DEVICE row offsets of the graph and of its transpose ([n + 1] each);
// results are device arrays of k ints
template <typename SizeT>
inline cudaError_t TopK(const SizeT *out_offsets, const SizeT *in_offsets, long long n, int k,
                        int *d_node_ids, int *d_in_degrees, int *d_out_degrees, int *launches, cudaStream_t s)
{
    if (n <= 0 || k <= 0) return cudaSuccess;
    util::DeviceArray<unsigned long long> a, b;
    util::RadixSortScratch scratch;
    GR_TRY(a.Allocate((size_t)n));
    GR_TRY(b.Allocate((size_t)n));
    DegreeKeysKernel<SizeT><<<util::FillGrid((unsigned long long)n), 256, 0, s>>>(out_offsets, in_offsets, n, a.ptr);
    GR_TRY(cudaGetLastError());
    unsigned long long *sorted = nullptr;
    GR_TRY(util::RadixSort64(a.ptr, b.ptr, (unsigned long long)n, 64, scratch, &sorted, s));
    GatherTopKernel<SizeT><<<(k + 255) / 256, 256, 0, s>>>(sorted, out_offsets, in_offsets, k,
                                                          d_node_ids, d_in_degrees, d_out_degrees);
    GR_TRY(cudaGetLastError());
    if (launches) *launches = 2 + 8 * 3;
    return cudaStreamSynchronize(s);        // a / b / scratch are released on return
}

}  // namespace topk
}  // namespace app
}  // namespace gunrock
