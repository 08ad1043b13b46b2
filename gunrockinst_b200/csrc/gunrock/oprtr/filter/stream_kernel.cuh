// stream_kernel.cuh -- the streaming instance of the filter operator for idempotent problems: big
// frontiers whose only cull is the exact claim of the vertex's bit in the visited mask.
//
// Same operator as filter::Kernel (kernel.cuh; reference: gunrock/oprtr/filter/cta.cuh:166-306,
// 467-544 -- BitmaskCull with a texture fetch per item, VertexCull, history / warp hash, raking
// scan, one atomic per 512-item tile), rebuilt on the machinery of advance/lb_stream.cuh, whose
// kernel it instantiates with DIRECT = true:
//   * the input frontier arrives in shared memory by one TMA 1-D bulk copy per 6144-item tile
//     (cp.async.bulk, three tiles in flight per SM, one mbarrier per ring slot);
//   * consumer warps probe the visited words of the hot vertices in a shared-memory copy and the
//     others at L2, eight items per lane in flight, and claim with fire-and-forget reductions --
//     duplicates of an item, the reason the reference needs its history and warp hashes, collapse
//     in the bit;
//   * the survivors are read off the mask afterwards (lb::HarvestKernel: set now, clear in the
//     snapshot taken before the launch): Functor::ApplyFilter and the compacted output frontier
//     are coalesced passes in vertex order, each vertex exactly once.
// A first TMA-fed filter of this round kept the generic kernel's exact claim (atomicOr, wait for
// the answer, label, staged enqueue) and was slower than filter::Kernel (2.85 against 2.21 ms on
// 274 M items at scale 25): the claim chain, not the input stream, bounds a filter.
//
// Restrictions (LaunchKernel of kernel.cuh is the general path): Problem::ENABLE_IDEMPOTENCE without
// predecessors, an explicit input queue, no predecessor queue, compaction, a single device, and a
// functor whose CondFilter only rejects holes.
#pragma once
#include "kernel.cuh"
#include "../advance/kernel.cuh"

namespace gunrock {
namespace oprtr {
namespace filter {

// callers check this before LaunchStreamKernel
template <typename Problem>
inline bool StreamEligible(const Args<typename Problem::VertexId, typename Problem::Value> &args, int log_parts)
{
    return Problem::ENABLE_IDEMPOTENCE && !Problem::MARK_PREDECESSORS && args.in_queue != nullptr &&
           args.in_pred_queue == nullptr && args.compact && log_parts == 0;
}

// AdvPolicy: an advance::KernelPolicy with THREADS = 1024, PRE_CULL on (the ring / consumer shape of lb_stream.cuh)
template <typename AdvPolicy, typename Problem, typename Functor>
cudaError_t LaunchStreamKernel(const advance::DeviceInfo &dev, const typename Problem::DataSlice &ds,
                               const Args<typename Problem::VertexId, typename Problem::Value> &fa,
                               const graph::CsrView<typename Problem::VertexId, typename Problem::SizeT> &g,
                               advance::LbScratch<typename Problem::SizeT> *scratch, unsigned hot_words, cudaStream_t stream)
{
    typedef typename Problem::VertexId VertexId;
    typedef typename Problem::SizeT    SizeT;
    namespace lb = advance::lb;
    auto kernel = lb::StreamExpandKernel<AdvPolicy, Problem, Functor, true>;
    static bool configured = false;     // per instantiation
    if (!configured) {
        GR_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
        configured = true;
    }
    if (hot_words > lb::StreamMaxHotWords<AdvPolicy>()) hot_words = lb::StreamMaxHotWords<AdvPolicy>();
    advance::Args<VertexId, SizeT> a;
    a.g = g;
    a.in_queue = fa.in_queue; a.out_queue = fa.out_queue;
    a.in_len = fa.in_len; a.d_in_len = fa.d_in_len; a.out_cap = fa.out_cap;
    a.slots = fa.slots; a.queue_index = fa.queue_index; a.iteration = fa.iteration;
    a.output = fa.out_queue ? advance::OUT_COMPACT : advance::OUT_NONE;
    advance::LbScratchView<SizeT> sv = {};
    // the visited mask as it is before the launch: the harvest reads the survivors off the difference
    const unsigned long long words = (unsigned long long)((g.nodes + 31) / 32);
    GR_TRY(scratch->mask_before.Allocate((size_t)words + 4));
    GR_TRY(cudaMemcpyAsync(scratch->mask_before.ptr, ds.d_visited_mask, sizeof(unsigned) * (size_t)words, cudaMemcpyDeviceToDevice, stream));
    unsigned long long tiles = (fa.in_len + AdvPolicy::TILE_EDGES - 1) / AdvPolicy::TILE_EDGES;
    unsigned grid = (unsigned)(tiles < 1 ? 1 : (tiles < (unsigned long long)dev.sm_count ? tiles : (unsigned long long)dev.sm_count));
    const size_t smem = lb::StreamHotOffset<AdvPolicy>() + sizeof(unsigned) * (size_t)((hot_words + 3u) & ~3u);
    kernel<<<grid, AdvPolicy::THREADS, smem, stream>>>(ds, a, sv, hot_words);
    GR_TRY(cudaGetLastError());
    unsigned long long hctas = (words + 255) / 256;
    const unsigned long long hmax = (unsigned long long)dev.sm_count * 8ull;
    lb::HarvestKernel<AdvPolicy, Problem, Functor, true><<<(unsigned)(hctas < 1 ? 1 : (hctas < hmax ? hctas : hmax)), 256, 0, stream>>>(
        ds, a, scratch->mask_before.ptr, words);
    return util::GRError(cudaGetLastError(), "filter::StreamKernel launch failed", __FILE__, __LINE__);
}

}  // namespace filter
}  // namespace oprtr
}  // namespace gunrock
