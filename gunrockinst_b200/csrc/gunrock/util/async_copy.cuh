// async_copy.cuh -- sm_100a asynchronous global -> shared copies and the mbarrier that tracks them.
//
// The reference moves every neighbour id through registers with (texture) loads issued by the
// thread that consumes it (gunrock/oprtr/edge_map_forward/cta.cuh:180-541,
// oprtr/edge_map_partitioned/kernel.cuh:242-559).  The streaming advance of lb_stream.cuh lets
// the copy engines do that instead: long runs of a neighbour list arrive in shared memory
// through TMA 1-D bulk copies (cp.async.bulk, SASS UBLKCP), the few unaligned entries around
// them and short lists through 4-byte cp.async (SASS LDGSTS), and one mbarrier per ring slot
// tells the consumers when a tile is complete.  Thin wrappers over the PTX, nothing else.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gunrock {
namespace util {

__device__ __forceinline__ unsigned SharedAddr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

// ---- mbarrier ---------------------------------------------------------------------------------------
__device__ __forceinline__ void MbarInit(unsigned long long *bar, unsigned arrivals)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(SharedAddr(bar)), "r"(arrivals) : "memory");
}
// make the initialised barriers visible to the async proxy (TMA completes on them)
__device__ __forceinline__ void MbarInitFence()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// add `bytes` to the transaction count the current phase waits for (no arrival)
__device__ __forceinline__ void MbarExpectTx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" :: "r"(SharedAddr(bar)), "r"(bytes) : "memory");
}
// one arrival that also adds `bytes` to the transaction count, as ONE operation on the barrier word: the form to use when
// the arrival follows the announcement immediately (an expect_tx and a cp.async-tracked arrival travel through different
// pipes and may reach the barrier in either order)
__device__ __forceinline__ void MbarArriveExpectTx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(SharedAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void MbarArrive(unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(SharedAddr(bar)) : "memory");
}
// one arrival that fires once every cp.async this thread has issued so far has landed (immediately
// when there is none); .noinc: it is one of the arrivals the barrier was initialised with
__device__ __forceinline__ void MbarArriveOnCpAsync(unsigned long long *bar)
{
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(SharedAddr(bar)) : "memory");
}
// as MbarArriveOnCpAsync, for a thread that is NOT one of the arrivals the barrier was initialised with: the pending count
// goes up by one now and down by one when the thread's cp.asyncs have landed, so the phase cannot complete before that
__device__ __forceinline__ void MbarJoinOnCpAsync(unsigned long long *bar)
{
    asm volatile("cp.async.mbarrier.arrive.shared::cta.b64 [%0];" :: "r"(SharedAddr(bar)) : "memory");
}
__device__ __forceinline__ bool MbarTryWait(unsigned long long *bar, unsigned parity)
{
    unsigned ok;
    asm volatile(
        "{\n .reg .pred p;\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        " selp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(SharedAddr(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void MbarWait(unsigned long long *bar, unsigned parity)
{
    while (!MbarTryWait(bar, parity)) { }
}

// ---- copies -------------------------------------------------------------------------------------------
// TMA 1-D bulk copy: `bytes` (multiple of 16) from 16-byte aligned global memory to 16-byte aligned
// shared memory; completion is reported to `bar` as `bytes` of its transaction count
__device__ __forceinline__ void BulkCopyG2S(void *smem_dst, const void *gmem_src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(SharedAddr(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(SharedAddr(bar)) : "memory");
}
// one 4-byte word, global -> shared, no register staging (Ampere-style cp.async, SASS LDGSTS)
__device__ __forceinline__ void CpAsync4(void *smem_dst, const void *gmem_src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(SharedAddr(smem_dst)), "l"(gmem_src) : "memory");
}

}  // namespace util
}  // namespace gunrock
