// scan.cuh -- device-wide exclusive prefix sum over 64-bit counters.
//
// Replaces the reference's calls into moderngpu's Scan (e.g. gunrock/oprtr/
// advance/kernel.cuh:312-313, priority_queue/kernel.cuh:435-436) for the places
// where a stand-alone scan is still wanted (graph construction, radix sort
// histograms, priority-queue splits).  Three launches, no inter-CTA waiting:
//   1. every CTA reduces its tile                        -> block_sums[b]
//   2. ONE CTA scans block_sums in place (exclusive)
//   3. every CTA re-reads its tile, scans it and adds its base, in place
// The hot advance path does not use this (it has its own fused single-pass
// scan, oprtr/advance/lb.cuh).
#pragma once
#include "array.cuh"
#include "device_intrinsics.cuh"

namespace gunrock {
namespace util {

constexpr int kScanThreads = 256;
constexpr int kScanItems   = 8;
constexpr int kScanTile    = kScanThreads * kScanItems;

template <typename T>
__global__ void __launch_bounds__(kScanThreads)
ScanReduceKernel(const T *in, unsigned long long n, unsigned long long *block_sums)
{
    __shared__ unsigned long long s_warp[kScanThreads / 32 + 1];
    const unsigned long long base = (unsigned long long)blockIdx.x * kScanTile;
    unsigned long long mine = 0;
#pragma unroll
    for (int r = 0; r < kScanItems; ++r) {
        const unsigned long long i = base + (unsigned long long)r * kScanThreads + threadIdx.x;
        if (i < n) mine += (unsigned long long)in[i];
    }
    unsigned long long total;
    (void)BlockExclusiveSum<kScanThreads>(mine, s_warp, total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// single CTA, in place; also writes the grand total to sums[count]
static __global__ void __launch_bounds__(1024)
ScanSpineKernel(unsigned long long *sums, unsigned long long count)
{
    __shared__ unsigned long long s_warp[1024 / 32 + 1];
    unsigned long long running = 0;
    for (unsigned long long base = 0; base < count; base += 1024) {
        const unsigned long long i = base + threadIdx.x;
        const unsigned long long v = (i < count) ? sums[i] : 0ull;
        unsigned long long total;
        const unsigned long long excl = BlockExclusiveSum<1024>(v, s_warp, total);
        if (i < count) sums[i] = running + excl;
        running += total;
    }
    if (threadIdx.x == 0) sums[count] = running;
}

// thread t owns kScanItems CONSECUTIVE elements (blocked arrangement) so the
// per-thread running sum is a plain loop
template <typename T>
__global__ void __launch_bounds__(kScanThreads)
ScanDownsweepKernel(const T *in, unsigned long long *out, unsigned long long n, const unsigned long long *block_sums)
{
    __shared__ unsigned long long s_warp[kScanThreads / 32 + 1];
    const unsigned long long base = (unsigned long long)blockIdx.x * kScanTile + (unsigned long long)threadIdx.x * kScanItems;
    unsigned long long v[kScanItems], mine = 0;
#pragma unroll
    for (int r = 0; r < kScanItems; ++r) {
        v[r] = (base + r < n) ? (unsigned long long)in[base + r] : 0ull;
        mine += v[r];
    }
    unsigned long long total;
    unsigned long long run = block_sums[blockIdx.x] + BlockExclusiveSum<kScanThreads>(mine, s_warp, total);
#pragma unroll
    for (int r = 0; r < kScanItems; ++r) {
        if (base + r < n) out[base + r] = run;
        run += v[r];
    }
}

// Exclusive scan of in[0..n) into out[0..n) (out may alias in when T is 64 bits wide);
// out_total (device, optional) receives the sum.  `block_sums` must hold n / kScanTile + 2.
template <typename T>
inline cudaError_t ExclusiveScan(const T *in, unsigned long long *out, unsigned long long n,
                                 unsigned long long *block_sums, cudaStream_t s)
{
    if (n == 0) return cudaMemsetAsync(block_sums, 0, sizeof(unsigned long long), s);
    const unsigned long long blocks = (n + kScanTile - 1) / kScanTile;
    ScanReduceKernel<T><<<(unsigned)blocks, kScanThreads, 0, s>>>(in, n, block_sums);
    GR_TRY(cudaGetLastError());
    ScanSpineKernel<<<1, 1024, 0, s>>>(block_sums, blocks);
    GR_TRY(cudaGetLastError());
    ScanDownsweepKernel<T><<<(unsigned)blocks, kScanThreads, 0, s>>>(in, out, n, block_sums);
    return GRError(cudaGetLastError(), "ExclusiveScan failed", __FILE__, __LINE__);
    // the grand total is left in block_sums[blocks]
}
inline unsigned long long ScanBlocks(unsigned long long n) { return (n + kScanTile - 1) / kScanTile; }

}  // namespace util
}  // namespace gunrock
