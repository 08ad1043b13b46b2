// radix_sort.cuh -- stable LSD radix sort of 64-bit keys on the device.
//
// Replaces the reference's dependency on CUB (util/sort_utils.cuh:32-120, the
// submodule is not even vendored) and on std::stable_sort in Csr::FromCoo
// (gunrock/csr.cuh:266-270) for graph construction.  8-bit digits; per pass:
//   1. HistKernel    per-tile digit histogram               -> hist[digit][tile]
//   2. ExclusiveScan over the digit-major histogram         (util/scan.cuh)
//   3. ScatterKernel re-reads the tile and moves every key to its final slot of
//      the pass.  Stability: a tile is cut into one contiguous chunk per warp,
//      a warp walks its chunk 32 keys at a time, and lanes holding the same
//      digit are ranked by lane with __match_any_sync.
// Graph construction is outside every timed region; the layout favours
// simplicity (no inter-CTA waiting) over the last 2x of sort throughput.
#pragma once
#include "scan.cuh"

namespace gunrock {
namespace util {

constexpr int kSortThreads = 256;
constexpr int kSortWarps   = kSortThreads / 32;
constexpr int kSortRows    = 32;                                  // 32-key rows per warp per tile
constexpr int kSortTile    = kSortWarps * 32 * kSortRows;         // 8192 keys per CTA
constexpr int kRadix       = 256;

static __global__ void __launch_bounds__(kSortThreads)
RadixHistKernel(const unsigned long long *keys, unsigned long long n, int shift,
                unsigned long long *hist, unsigned long long tiles)
{
    __shared__ unsigned s_h[kRadix];
    s_h[threadIdx.x] = 0;
    __syncthreads();
    const unsigned long long base = (unsigned long long)blockIdx.x * kSortTile;
#pragma unroll 4
    for (int r = 0; r < kSortTile / kSortThreads; ++r) {
        const unsigned long long i = base + (unsigned long long)r * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&s_h[(unsigned)(keys[i] >> shift) & (kRadix - 1)], 1u);
    }
    __syncthreads();
    hist[(unsigned long long)threadIdx.x * tiles + blockIdx.x] = s_h[threadIdx.x];
}

static __global__ void __launch_bounds__(kSortThreads)
RadixScatterKernel(const unsigned long long *in, unsigned long long *out, unsigned long long n, int shift,
                   const unsigned long long *scanned, unsigned long long tiles)
{
    __shared__ unsigned           s_count[kSortWarps][kRadix];
    __shared__ unsigned long long s_base[kSortWarps][kRadix];
    const int lane = LaneId(), warp = WarpId();
    for (int k = threadIdx.x; k < kSortWarps * kRadix; k += kSortThreads) (&s_count[0][0])[k] = 0;
    __syncthreads();

    const unsigned long long chunk = (unsigned long long)blockIdx.x * kSortTile + (unsigned long long)warp * 32 * kSortRows;
    // phase A: digit counts of this warp's chunk
    for (int r = 0; r < kSortRows; ++r) {
        const unsigned long long i = chunk + (unsigned long long)r * 32 + lane;
        const bool live = i < n;
        const unsigned d = live ? ((unsigned)(in[i] >> shift) & (kRadix - 1)) : (unsigned)(kRadix + lane);
        const unsigned peers = __match_any_sync(kFullMask, d);
        if (live && (__ffs(peers) - 1) == lane) s_count[warp][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    {   // thread d: where each warp's run of digit d starts
        const int d = threadIdx.x;
        unsigned long long run = scanned[(unsigned long long)d * tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) { s_base[w][d] = run; run += s_count[w][d]; }
    }
    __syncthreads();
    // phase B: rank and move
    for (int r = 0; r < kSortRows; ++r) {
        const unsigned long long i = chunk + (unsigned long long)r * 32 + lane;
        const bool live = i < n;
        unsigned long long key = 0;
        if (live) key = in[i];
        const unsigned d = live ? ((unsigned)(key >> shift) & (kRadix - 1)) : (unsigned)(kRadix + lane);
        const unsigned peers = __match_any_sync(kFullMask, d);
        if (live) {
            const unsigned long long b = s_base[warp][d];
            out[b + __popc(peers & ((1u << lane) - 1))] = key;
        }
        __syncwarp();
        if (live && (__ffs(peers) - 1) == lane) s_base[warp][d] += __popc(peers);
        __syncwarp();
    }
}

struct RadixSortScratch {
    DeviceArray<unsigned long long> hist;        // [kRadix * tiles]
    DeviceArray<unsigned long long> block_sums;
    cudaError_t Allocate(unsigned long long n)
    {
        const unsigned long long tiles = (n + kSortTile - 1) / kSortTile;
        GR_TRY(hist.Allocate((size_t)(kRadix * (tiles ? tiles : 1))));
        return block_sums.Allocate((size_t)(ScanBlocks(kRadix * (tiles ? tiles : 1)) + 2));
    }
};

// Sorts the low `key_bits` bits ascending.  On return *sorted points at whichever of
// a / b holds the result.
inline cudaError_t RadixSort64(unsigned long long *a, unsigned long long *b, unsigned long long n, int key_bits,
                               RadixSortScratch &scratch, unsigned long long **sorted, cudaStream_t s)
{
    *sorted = a;
    if (n < 2) return cudaSuccess;
    GR_TRY(scratch.Allocate(n));
    const unsigned long long tiles = (n + kSortTile - 1) / kSortTile;
    unsigned long long *src = a, *dst = b;
    for (int shift = 0; shift < key_bits; shift += 8) {
        RadixHistKernel<<<(unsigned)tiles, kSortThreads, 0, s>>>(src, n, shift, scratch.hist.ptr, tiles);
        GR_TRY(cudaGetLastError());
        GR_TRY(ExclusiveScan<unsigned long long>(scratch.hist.ptr, scratch.hist.ptr, kRadix * tiles,
                                                 scratch.block_sums.ptr, s));
        RadixScatterKernel<<<(unsigned)tiles, kSortThreads, 0, s>>>(src, dst, n, shift, scratch.hist.ptr, tiles);
        GR_TRY(cudaGetLastError());
        unsigned long long *t = src; src = dst; dst = t;
    }
    *sorted = src;
    return cudaSuccess;
}

}  // namespace util
}  // namespace gunrock
