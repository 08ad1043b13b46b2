// row_bins.cuh -- a fixed set of CSR rows binned by length, for passes that walk the same rows
// every time (every sweep of PageRank's power iteration, the hook pass of connected components).
//
// The reference hands such passes to its load-balanced advance like any other frontier
// (gunrock/oprtr/edge_map_partitioned/kernel.cuh:242-559: degree scan, sorted search, one thread
// per entry) and so does oprtr/advance/lb.cuh here; for a row set that is known before the pass
// starts, the scan and the per-entry owner search buy nothing.  The rows go into three lists,
//     small   1 .. 16 entries       four lanes per row (eight rows per warp),
//     medium  17 .. 2048            one warp per row,
//     big     > 2048                cut into 2048-entry pieces, one warp per piece,
// and one launch covers them all: with a static grid CTAs [0, blocks_big) take the pieces (the
// longest work first), the next blocks_medium CTAs the medium rows, the rest the small rows; with a
// resident grid the warps draw pieces, then groups of medium rows, then groups of small rows from a
// ticket counter (pr_row_sweep.cuh).  Rows without entries are in no list.
// Users: app/pr/pr_row_sweep.cuh, app/cc/cc_row_hook.cuh; stand-alone: grb_test_row_bins.
#pragma once
#include "../graph/csr.cuh"
#include "../util/array.cuh"
#include "../util/memset_kernel.cuh"
#include "../util/device_intrinsics.cuh"
#include "advance/common.cuh"

namespace gunrock {
namespace oprtr {
namespace rows {

constexpr int kSmallMax      = 16;      // longest row of the 4-lane bin
constexpr int kChunk         = 2048;    // longest row one warp takes whole; piece size of longer rows
constexpr int kSmallIters    = 4;       // rounds of 8 small rows per warp
constexpr int kMediumPerWarp = 4;       // medium rows per warp
constexpr int kThreads       = 256;     // CTA size of a pass over the bins
constexpr int kWarps         = kThreads / 32;

enum { SMALL = 0, MEDIUM = 1, BIG = 2, PIECES = 3, ENTRIES = 4, COUNTERS = 8 };

// what a kernel needs of the bins (by value in the parameter bank)
struct View {
    const int  *small_rows, *medium_rows, *big_rows;    // local row ids
    const int2 *big_pieces;                             // (row, piece index)
    unsigned long long n_small, n_medium, n_big, n_pieces, entries;
    unsigned blocks_big, blocks_medium, blocks_small;
    unsigned Blocks() const { return blocks_big + blocks_medium + blocks_small; }
};

// rows queue[0 .. qlen) (global ids; nullptr: local rows 0 .. qlen) -> the three lists (order follows
// the queue within a CTA's round of 256 rows) and the piece list; counts[ENTRIES] = entries of all listed rows
template <typename SizeT>
__global__ void __launch_bounds__(256)
BinKernel(const SizeT *row_offsets, const int *queue, unsigned long long qlen, graph::Partition part,
          int *small_rows, int *medium_rows, int *big_rows, int2 *big_pieces, unsigned long long *counts)
{
    __shared__ unsigned long long s_red[256 / 32 + 1];
    __shared__ unsigned s_cnt[256 / 32][3];             // per warp and bin: rows of this round, then their offset in the CTA's run
    __shared__ unsigned long long s_base[3];            // where the CTA's run of this round starts in each list
    const unsigned long long rounded = (qlen + 255) & ~255ull;      // whole CTAs go round together
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned lane = (unsigned)util::LaneId(), warp = (unsigned)util::WarpId();
    const unsigned below = (1u << lane) - 1u;
    unsigned long long entries = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < rounded; i += stride) {
        int r = -1;
        long long len = 0;
        if (i < qlen) {
            r = queue ? (int)part.Local(queue[i]) : (int)i;
            len = (long long)(row_offsets[r + 1] - row_offsets[r]);
        }
        entries += (unsigned long long)len;
        const int bin = len <= 0 ? -1 : (len <= kSmallMax ? SMALL : (len <= kChunk ? MEDIUM : BIG));
        // one atomic per CTA, round and list (one per warp serialised the pass on three words: 1.1 ms at scale 25)
        unsigned m[3];
#pragma unroll
        for (int b = 0; b < 3; ++b) m[b] = __ballot_sync(util::kFullMask, bin == b);
        if (lane == 0) {
#pragma unroll
            for (int b = 0; b < 3; ++b) s_cnt[warp][b] = (unsigned)__popc(m[b]);
        }
        __syncthreads();
        if (threadIdx.x < 3) {
            unsigned total = 0;
            for (int w = 0; w < 256 / 32; ++w) { const unsigned c = s_cnt[w][threadIdx.x]; s_cnt[w][threadIdx.x] = total; total += c; }
            s_base[threadIdx.x] = total ? atomicAdd(&counts[threadIdx.x], (unsigned long long)total) : 0ull;
        }
        __syncthreads();
        if (bin >= 0) {
            int *list = bin == SMALL ? small_rows : (bin == MEDIUM ? medium_rows : big_rows);
            list[s_base[bin] + s_cnt[warp][bin] + (unsigned)__popc(m[bin] & below)] = r;
        }
        __syncthreads();                                // s_cnt / s_base are rewritten by the next round
        if (bin == BIG) {
            const unsigned long long pieces = (unsigned long long)((len + kChunk - 1) / kChunk);
            const unsigned long long at = atomicAdd(&counts[PIECES], pieces);
            for (unsigned long long k = 0; k < pieces; ++k) big_pieces[at + k] = make_int2(r, (int)k);
        }
    }
    advance::BlockAccumulate<256>(entries, s_red, &counts[ENTRIES]);
}

// owner of the lists
template <typename SizeT>
struct Bins {
    util::DeviceArray<int>  small_rows, medium_rows, big_rows;
    util::DeviceArray<int2> big_pieces;
    util::DeviceArray<unsigned long long> counts;
    util::PinnedArray<unsigned long long> h_counts;
    View view;

    // room for any subset of a graph slice with `rows` rows and `entries` entries
    cudaError_t Allocate(size_t rows, size_t entries)
    {
        const size_t med = entries / (kSmallMax + 1);
        GR_TRY(small_rows.Allocate(rows + 1));
        GR_TRY(medium_rows.Allocate((med < rows ? med : rows) + 1));
        GR_TRY(big_rows.Allocate(entries / (kChunk + 1) + 1));
        GR_TRY(big_pieces.Allocate(2 * (entries / kChunk) + 2));
        GR_TRY(counts.Allocate(COUNTERS));
        return h_counts.Allocate(COUNTERS);
    }

    // one kernel and one blocking 64-byte read-back; fills `view`
    cudaError_t Build(const SizeT *row_offsets, const int *queue, unsigned long long qlen, graph::Partition part,
                      cudaStream_t stream)
    {
        GR_TRY(cudaMemsetAsync(counts.ptr, 0, sizeof(unsigned long long) * COUNTERS, stream));
        BinKernel<SizeT><<<util::FillGrid(qlen + 32), 256, 0, stream>>>(
            row_offsets, queue, qlen, part, small_rows.ptr, medium_rows.ptr, big_rows.ptr, big_pieces.ptr, counts.ptr);
        GR_TRY(cudaGetLastError());
        GR_TRY(cudaMemcpyAsync(h_counts.ptr, counts.ptr, sizeof(unsigned long long) * COUNTERS, cudaMemcpyDeviceToHost, stream));
        GR_TRY(cudaStreamSynchronize(stream));
        const unsigned long long *h = h_counts.ptr;
        View v;
        v.small_rows = small_rows.ptr; v.medium_rows = medium_rows.ptr; v.big_rows = big_rows.ptr; v.big_pieces = big_pieces.ptr;
        v.n_small = h[SMALL]; v.n_medium = h[MEDIUM]; v.n_big = h[BIG]; v.n_pieces = h[PIECES]; v.entries = h[ENTRIES];
        if (v.n_small + v.n_medium + v.n_big > qlen || v.n_small > small_rows.cap || v.n_medium > medium_rows.cap ||
            v.n_big > big_rows.cap || v.n_pieces > big_pieces.cap)
            return util::GRError(cudaErrorUnknown, "row bins do not add up", __FILE__, __LINE__);
        const unsigned long long w = kWarps;
        v.blocks_big    = (unsigned)((v.n_pieces + w - 1) / w);
        v.blocks_medium = (unsigned)((v.n_medium + w * kMediumPerWarp - 1) / (w * kMediumPerWarp));
        v.blocks_small  = (unsigned)((v.n_small + w * 8 * kSmallIters - 1) / (w * 8 * kSmallIters));
        view = v;
        return cudaSuccess;
    }
};

// entries [beg, end) of piece `piece` of the big rows
template <typename SizeT>
__device__ __forceinline__ int PieceRange(const View &b, const SizeT *row_offsets, unsigned long long piece, SizeT &beg, SizeT &end)
{
    const int2 it = b.big_pieces[piece];
    const SizeT row_end = row_offsets[it.x + 1];
    beg = row_offsets[it.x] + (SizeT)it.y * (SizeT)kChunk;
    end = (row_end - beg > (SizeT)kChunk) ? beg + (SizeT)kChunk : row_end;
    return it.x;
}

}  // namespace rows
}  // namespace oprtr
}  // namespace gunrock
