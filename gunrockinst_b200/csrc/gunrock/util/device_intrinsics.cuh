// device_intrinsics.cuh -- sm_100a load/store flavours and warp-level helpers.
//
// Replaces the reference's util/io/modified_load.cuh / modified_store.cuh (PTX
// cache-modifier wrappers written for sm_20/sm_35) and its smem raking scans
// (util/scan/cooperative_scan.cuh) with warp shuffle / ballot primitives.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gunrock {
namespace util {

constexpr int kWarpThreads = 32;
constexpr unsigned kFullMask = 0xffffffffu;

__device__ __forceinline__ int LaneId() { return threadIdx.x & 31; }
__device__ __forceinline__ int WarpId() { return threadIdx.x >> 5; }

// ---- loads ------------------------------------------------------------------
// streaming, read-only data that is touched once (column indices, queues):
// non-coherent path, do not keep in L1
template <typename T> __device__ __forceinline__ T LoadStream(const T *p);
template <> __device__ __forceinline__ int LoadStream<int>(const int *p)
{
    int v; asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p)); return v;
}
template <> __device__ __forceinline__ unsigned LoadStream<unsigned>(const unsigned *p)
{
    unsigned v; asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p)); return v;
}
template <> __device__ __forceinline__ long long LoadStream<long long>(const long long *p)
{
    long long v; asm volatile("ld.global.nc.L1::no_allocate.s64 %0, [%1];" : "=l"(v) : "l"(p)); return v;
}
template <> __device__ __forceinline__ float LoadStream<float>(const float *p)
{
    float v; asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p)); return v;
}

// four consecutive ints in one request (p must be 16-byte aligned)
__device__ __forceinline__ int4 LoadStream4(const int *p)
{
    int4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// data another SM may be updating in this same launch (labels, bitmaps, ranks):
// read at L2, never from a stale L1 line
template <typename T> __device__ __forceinline__ T LoadCG(const T *p) { return __ldcg(p); }
// read-mostly probe where a stale L1 hit is harmless (a later atomic decides)
template <typename T> __device__ __forceinline__ T LoadCA(const T *p) { return __ldca(p); }
// write-through store that does not pollute L1
template <typename T> __device__ __forceinline__ void StoreCG(T *p, T v) { __stcg(p, v); }

// ---- warp reductions / scans ---------------------------------------------------
template <typename T>
__device__ __forceinline__ T WarpReduceSum(T v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}

// inclusive prefix sum across the warp
template <typename T>
__device__ __forceinline__ T WarpInclusiveSum(T v)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T t = __shfl_up_sync(kFullMask, v, o);
        if (LaneId() >= o) v += t;
    }
    return v;
}

// Exclusive prefix sum over a CTA of THREADS threads (THREADS % 32 == 0,
// THREADS <= 1024).  `warp_totals` is THREADS/32 + 1 words of shared memory.
// Returns the exclusive prefix of `v`; `total` receives the CTA-wide sum.
// Contains two __syncthreads(): call from convergent code only.
template <int THREADS, typename T>
__device__ __forceinline__ T BlockExclusiveSum(T v, T *warp_totals, T &total)
{
    constexpr int WARPS = THREADS / 32;
    T incl = WarpInclusiveSum(v);
    if (LaneId() == 31) warp_totals[WarpId()] = incl;
    __syncthreads();
    if (WarpId() == 0) {
        T w = (LaneId() < WARPS) ? warp_totals[LaneId()] : T(0);
        T wi = WarpInclusiveSum(w);
        if (LaneId() < WARPS) warp_totals[LaneId()] = wi - w;
        if (LaneId() == 31) warp_totals[WARPS] = wi;
    }
    __syncthreads();
    T base = warp_totals[WarpId()];
    total = warp_totals[WARPS];
    T r = base + incl - v;
    __syncthreads();            // warp_totals may be reused immediately by the caller
    return r;
}

// ---- bitmap helpers --------------------------------------------------------------
__device__ __forceinline__ bool BitTest(const unsigned *bitmap, unsigned long long pos)
{
    return (LoadCA(bitmap + (pos >> 5)) >> (pos & 31)) & 1u;
}
// set the bit; true iff THIS call changed it (unique winner among racing callers)
__device__ __forceinline__ bool BitTestAndSet(unsigned *bitmap, unsigned long long pos)
{
    unsigned bit = 1u << (pos & 31);
    unsigned old = atomicOr(bitmap + (pos >> 5), bit);
    return (old & bit) == 0;
}

// Predicated forms: the atomic is issued under a predicate instead of inside a branch, so that a thread
// that claims several bits in a row has all its atomics in flight at once (inside `if` bodies the
// compiler tests every answer before it issues the next atomic: one L2 round trip per claim).
// AtomicOrIf: the word before the OR, or `bit` itself (= "was already set") when `p` is false.
__device__ __forceinline__ unsigned AtomicOrIf(bool p, unsigned *addr, unsigned bit)
{
    unsigned old = bit;
    asm volatile("{\n .reg .pred q;\n setp.ne.u32 q, %3, 0;\n @q atom.global.or.b32 %0, [%1], %2;\n}\n"
                 : "+r"(old) : "l"(addr), "r"(bit), "r"((unsigned)p) : "memory");
    return old;
}
// true iff `p` and THIS call changed the bit
__device__ __forceinline__ bool BitTestAndSetIf(bool p, unsigned *bitmap, unsigned long long pos)
{
    const unsigned bit = 1u << (pos & 31);
    return (AtomicOrIf(p, bitmap + (pos >> 5), bit) & bit) == 0;
}

}  // namespace util
}  // namespace gunrock
