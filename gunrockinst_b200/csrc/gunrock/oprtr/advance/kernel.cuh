// kernel.cuh -- advance::LaunchKernel, the host-side mode dispatcher.
//
// Counterpart of the reference's advance::LaunchKernel<KernelPolicy, Problem,
// Functor>(...) (gunrock/oprtr/advance/kernel.cuh:101-765): same template
// triple, same four modes, but every mode is launched without a host round
// trip -- the reference's LB branch blocks on a cudaMemcpy of the scanned total
// (:315-317) and, in this fork, on an RPC to an external scheduler daemon plus
// a relaunch loop of an instrumented kernel (:552-726); neither exists here.
#pragma once
#include "lb.cuh"
#include "lb_stream.cuh"
#include "twc.cuh"
#include "pull.cuh"
#include "reduce.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {

// what the launcher needs to know about the device
struct DeviceInfo {
    int device   = 0;
    int sm_count = 148;
    cudaError_t Query(int dev)
    {
        device = dev;
        // one attribute, not cudaGetDeviceProperties (which costs ~1 ms and sat in every one-shot call)
        return cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
    }
};

// Push-direction advance (TWC_FORWARD or LB).
//   edges_hint: upper bound on (or exact value of) the number of neighbour entries
//   the frontier expands to; it only sizes the persistent grid.
template <typename Policy, typename Problem, typename Functor>
cudaError_t LaunchKernel(
    MODE mode,
    const DeviceInfo &dev,
    const typename Problem::DataSlice &ds,
    Args<typename Policy::VertexId, typename Policy::SizeT> args,
    LbScratch<typename Policy::SizeT> *scratch,
    unsigned long long edges_hint,
    cudaStream_t stream)
{
    const unsigned long long resident = (unsigned long long)dev.sm_count * Policy::MIN_CTAS;
    if (mode == TWC_FORWARD) {
        unsigned long long tiles = (args.in_len + Policy::THREADS - 1) / Policy::THREADS;
        unsigned grid = (unsigned)(tiles < 1 ? 1 : (tiles < resident ? tiles : resident));
        twc::Kernel<Policy, Problem, Functor><<<grid, Policy::THREADS, 0, stream>>>(ds, args);
        return util::GRError(cudaGetLastError(), "advance twc::Kernel launch failed", __FILE__, __LINE__);
    }
    if (mode == LB) {
        if (!scratch) return util::GRError(cudaErrorInvalidValue, "advance LB needs scratch", __FILE__, __LINE__);
        if (args.in_len + 1 > scratch->scanned.cap)
            return util::GRError(cudaErrorInvalidValue, "advance LB scratch too small", __FILE__, __LINE__);
        if (++scratch->epoch >= lb::kEpochLimit) {          // descriptor tags wrapped: start over
            GR_TRY(scratch->tile_desc.Zero(stream));
            scratch->epoch = 1;
        }
        LbScratchView<typename Policy::SizeT> sv;
        sv.scanned = scratch->scanned.ptr;   sv.row_start = scratch->row_start.ptr;
        sv.tile_desc = scratch->tile_desc.ptr; sv.part_start = scratch->part_start.ptr; sv.ticket = scratch->ticket.ptr;
        sv.part_cap = scratch->part_start.cap;
        sv.epoch = scratch->epoch;
        unsigned long long scan_tiles = (args.in_len + Policy::SCAN_TILE - 1) / Policy::SCAN_TILE;
        unsigned scan_grid = (unsigned)(scan_tiles < 1 ? 1 : scan_tiles);
        lb::ScanKernel<Policy><<<scan_grid, Policy::SCAN_THREADS, 0, stream>>>(args, sv);
        GR_TRY(cudaGetLastError());
        unsigned long long tiles = (edges_hint + Policy::TILE_EDGES - 1) / Policy::TILE_EDGES;
        unsigned grid = (unsigned)(tiles < 1 ? 1 : (tiles < resident ? tiles : resident));
        lb::ExpandKernel<Policy, Problem, Functor><<<grid, Policy::THREADS, 0, stream>>>(ds, args, sv);
        return util::GRError(cudaGetLastError(), "advance lb::ExpandKernel launch failed", __FILE__, __LINE__);
    }
    return util::GRError(cudaErrorInvalidValue, "advance: push launcher called with a pull mode", __FILE__, __LINE__);
}

// The streaming instance of the LB mode (lb_stream.cuh): same scan, then one 1024-thread CTA per SM
// whose producer warps stage dense neighbour-id tiles in shared memory through TMA / cp.async and whose
// consumer warps probe the hot part of the visited map there.  Policy::THREADS must be 1024; callers check StreamEligible first.
template <typename Policy, typename Problem>
inline bool StreamEligible(const Args<typename Policy::VertexId, typename Policy::SizeT> &args)
{
    return !Problem::MARK_PREDECESSORS && !Problem::ENABLE_IDEMPOTENCE && Policy::PRE_CULL && args.type == V2V &&
           args.output != OUT_TOMBSTONE && args.g.part.log_parts == 0 && args.in_queue != nullptr;
}

template <typename Policy, typename Problem, typename Functor>
cudaError_t LaunchStreamKernel(
    const DeviceInfo &dev,
    const typename Problem::DataSlice &ds,
    Args<typename Policy::VertexId, typename Policy::SizeT> args,
    LbScratch<typename Policy::SizeT> *scratch,
    unsigned long long edges_hint,
    unsigned hot_words,                 // words of the visited map mirrored in shared memory (PRE_CULL policies)
    cudaStream_t stream)
{
    if (!scratch) return util::GRError(cudaErrorInvalidValue, "advance LB needs scratch", __FILE__, __LINE__);
    if (args.in_len + 1 > scratch->scanned.cap)
        return util::GRError(cudaErrorInvalidValue, "advance LB scratch too small", __FILE__, __LINE__);
    auto kernel = lb::StreamExpandKernel<Policy, Problem, Functor>;
    static bool configured = false;     // per instantiation
    if (!configured) {
        GR_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
        configured = true;
    }
    if (!Policy::PRE_CULL) hot_words = 0;
    if (hot_words > lb::StreamMaxHotWords<Policy>()) hot_words = lb::StreamMaxHotWords<Policy>();
    if (++scratch->epoch >= lb::kEpochLimit) {
        GR_TRY(scratch->tile_desc.Zero(stream));
        scratch->epoch = 1;
    }
    LbScratchView<typename Policy::SizeT> sv;
    sv.scanned = scratch->scanned.ptr;   sv.row_start = scratch->row_start.ptr;
    sv.tile_desc = scratch->tile_desc.ptr; sv.part_start = scratch->part_start.ptr; sv.ticket = scratch->ticket.ptr;
    sv.part_cap = scratch->part_start.cap;
    sv.epoch = scratch->epoch;
    unsigned long long scan_tiles = (args.in_len + Policy::SCAN_TILE - 1) / Policy::SCAN_TILE;
    unsigned scan_grid = (unsigned)(scan_tiles < 1 ? 1 : scan_tiles);
    lb::ScanKernel<Policy><<<scan_grid, Policy::SCAN_THREADS, 0, stream>>>(args, sv);
    GR_TRY(cudaGetLastError());
    // the visited mask as it is before the level: HarvestKernel reads the winners off the difference
    const unsigned long long words = (unsigned long long)((args.g.nodes + 31) / 32);
    GR_TRY(scratch->mask_before.Allocate((size_t)words + 4));
    GR_TRY(cudaMemcpyAsync(scratch->mask_before.ptr, ds.d_visited_mask, sizeof(unsigned) * (size_t)words, cudaMemcpyDeviceToDevice, stream));
    // one CTA per SM, each taking every gridDim.x-th tile
    unsigned long long tiles = (edges_hint + Policy::TILE_EDGES - 1) / Policy::TILE_EDGES;
    unsigned grid = (unsigned)(tiles < 1 ? 1 : (tiles < (unsigned long long)dev.sm_count ? tiles : (unsigned long long)dev.sm_count));
    const size_t smem = lb::StreamHotOffset<Policy>() + sizeof(unsigned) * (size_t)((hot_words + 3u) & ~3u);
    kernel<<<grid, Policy::THREADS, smem, stream>>>(ds, args, sv, hot_words);
    GR_TRY(cudaGetLastError());
    unsigned long long hctas = (words + 255) / 256;                          // a warp per 32 words, 8 warps per CTA
    const unsigned long long hmax = (unsigned long long)dev.sm_count * 8ull;
    lb::HarvestKernel<Policy, Problem, Functor><<<(unsigned)(hctas < 1 ? 1 : (hctas < hmax ? hctas : hmax)), 256, 0, stream>>>(
        ds, args, scratch->mask_before.ptr, words);
    return util::GRError(cudaGetLastError(), "advance lb::StreamExpandKernel launch failed", __FILE__, __LINE__);
}

// Pull-direction advance (TWC_BACKWARD / LB_BACKWARD share one implementation).
template <typename Policy, typename Problem, typename Functor>
cudaError_t LaunchPullKernel(
    const DeviceInfo &dev,
    const typename Problem::DataSlice &ds,
    pull::PullArgs<typename Policy::VertexId, typename Policy::SizeT> args,
    cudaStream_t stream)
{
    const unsigned long long resident = (unsigned long long)dev.sm_count * Policy::MIN_CTAS;
    unsigned long long words = (unsigned long long)((args.g.local_nodes + 31) >> 5);
    unsigned long long chunks = (words + pull::kChunkWords - 1) / pull::kChunkWords;
    unsigned long long ctas = (chunks + Policy::WARPS - 1) / Policy::WARPS;
    unsigned grid = (unsigned)(ctas < 1 ? 1 : (ctas < resident ? ctas : resident));
    pull::Kernel<Policy, Problem, Functor><<<grid, Policy::THREADS, 0, stream>>>(ds, args);
    return util::GRError(cudaGetLastError(), "advance pull::Kernel launch failed", __FILE__, __LINE__);
}

}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
