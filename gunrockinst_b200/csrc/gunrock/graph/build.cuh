// build.cuh -- device-side graph construction: sorted (row, col) tuples -> CSR, the transpose
// (CSC) of a resident CSR, and the degree-ordered twin the BFS kernels run on.
//
// Counterparts in the reference:
//   Csr::FromCoo                      gunrock/csr.cuh:247-340      (host: std::stable_sort + dedup)
//   the inverse graph of DOBFS        tests/dobfs/test_dobfs.cu:477-560 (a second, column-major
//                                     Csr built on the host and handed to DOBFSProblem::Init,
//                                     app/dobfs/dobfs_problem.cuh:211-260)
// The reference has nothing like the degree-ordered twin; see DESIGN.md section 2 for why a B200
// wants it (a 1-bit-per-vertex map of the ~1.5 M highest-degree vertices fits one SM's shared
// memory and answers 80-95 % of all visited / frontier probes of an R-MAT search there).
//
// Everything here is one-time construction outside any timed region; keys are 64-bit
// (row << bits | col) sorted by util::RadixSort64.
#pragma once
#include "csr.cuh"
#include "../util/radix_sort.cuh"
#include "../util/memset_kernel.cuh"

namespace gunrock {
namespace graph {

constexpr int kKeepThreads = 256, kKeepItems = 8, kKeepTile = kKeepThreads * kKeepItems;

// canonical: drop duplicates and self loops (csr.cuh:282-288); otherwise keep every tuple
__device__ __forceinline__ bool KeepTuple(const unsigned long long *keys, unsigned long long i, int bits,
                                          const Partition &part, bool canonical)
{
    if (!canonical) return true;
    const unsigned long long k = keys[i];
    if (i > 0 && keys[i - 1] == k) return false;
    const long long row = part.Global((long long)(k >> bits));
    const long long col = (long long)(k & ((1ull << bits) - 1));
    return row != col;
}

static __global__ void __launch_bounds__(kKeepThreads)
CountKeptKernel(const unsigned long long *keys, unsigned long long n, int bits, Partition part, bool canonical,
                unsigned long long *tile_counts)
{
    __shared__ unsigned long long s_warp[kKeepThreads / 32 + 1];
    const unsigned long long base = (unsigned long long)blockIdx.x * kKeepTile + (unsigned long long)threadIdx.x * kKeepItems;
    unsigned long long mine = 0;
#pragma unroll
    for (int r = 0; r < kKeepItems; ++r)
        if (base + r < n && KeepTuple(keys, base + r, bits, part, canonical)) ++mine;
    unsigned long long total;
    (void)util::BlockExclusiveSum<kKeepThreads>(mine, s_warp, total);
    if (threadIdx.x == 0) tile_counts[blockIdx.x] = total;
}

// writes col[j] and krow[j] (local row) for every kept tuple j
static __global__ void __launch_bounds__(kKeepThreads)
CompactKeptKernel(const unsigned long long *keys, unsigned long long n, int bits, Partition part, bool canonical,
                  const unsigned long long *tile_offsets, int *col, int *krow)
{
    __shared__ unsigned long long s_warp[kKeepThreads / 32 + 1];
    const unsigned long long base = (unsigned long long)blockIdx.x * kKeepTile + (unsigned long long)threadIdx.x * kKeepItems;
    bool keep[kKeepItems];
    unsigned long long mine = 0;
#pragma unroll
    for (int r = 0; r < kKeepItems; ++r) {
        keep[r] = base + r < n && KeepTuple(keys, base + r, bits, part, canonical);
        if (keep[r]) ++mine;
    }
    unsigned long long total;
    unsigned long long at = tile_offsets[blockIdx.x] + util::BlockExclusiveSum<kKeepThreads>(mine, s_warp, total);
#pragma unroll
    for (int r = 0; r < kKeepItems; ++r) {
        if (keep[r]) {
            const unsigned long long k = keys[base + r];
            col[at]  = (int)(k & ((1ull << bits) - 1));
            krow[at] = (int)(k >> bits);
            ++at;
        }
    }
}

// row_offsets[x] = first kept tuple whose local row is >= x
template <typename SizeT>
__global__ void RowOffsetsKernel(const int *krow, unsigned long long m, long long local_nodes, SizeT *row_offsets)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; j <= m; j += stride) {
        const long long lo = (j == 0) ? 0 : (long long)krow[j - 1] + 1;
        const long long hi = (j == m) ? local_nodes : (long long)krow[j];
        for (long long x = lo; x <= hi; ++x) row_offsets[x] = (SizeT)j;
    }
}

// Tuples sorted by key (local row << bits | col) -> row_offsets / col_indices.  `spare` is a
// buffer of at least `tuples` 64-bit words that is overwritten (it ends up holding the local row
// of every kept tuple as int32, returned through krow_out for callers that derive per-edge data).
// Returns cudaErrorInvalidValue when the slice needs 64-bit offsets and SizeT is 32 bits wide.
template <typename SizeT>
cudaError_t SortedKeysToCsr(const unsigned long long *sorted, unsigned long long *spare, unsigned long long tuples,
                            int bits, Partition part, bool canonical, long long local_nodes,
                            util::DeviceArray<SizeT> &row_offsets, util::DeviceArray<int> &col_indices,
                            unsigned long long *kept_out, int **krow_out, cudaStream_t s)
{
    const unsigned long long tiles = (tuples + kKeepTile - 1) / kKeepTile;
    util::DeviceArray<unsigned long long> tile_counts, block_sums;
    GR_TRY(tile_counts.Allocate((size_t)tiles + 2));
    GR_TRY(block_sums.Allocate((size_t)util::ScanBlocks(tiles + 1) + 2));
    unsigned long long m = 0;
    if (tiles) {
        CountKeptKernel<<<(unsigned)tiles, kKeepThreads, 0, s>>>(sorted, tuples, bits, part, canonical, tile_counts.ptr);
        GR_TRY(cudaGetLastError());
        GR_TRY(util::ExclusiveScan<unsigned long long>(tile_counts.ptr, tile_counts.ptr, tiles, block_sums.ptr, s));
        GR_TRY(cudaMemcpyAsync(&m, block_sums.ptr + util::ScanBlocks(tiles), sizeof(m), cudaMemcpyDeviceToHost, s));
        GR_TRY(cudaStreamSynchronize(s));
    }
    if (sizeof(SizeT) == 4 && m >= 0x7fffffffull)
        return util::GRError(cudaErrorInvalidValue, "graph slice needs 64-bit edge offsets", __FILE__, __LINE__);
    GR_TRY(col_indices.Allocate((size_t)m));
    GR_TRY(row_offsets.Allocate((size_t)local_nodes + 1));
    int *krow = reinterpret_cast<int *>(spare);
    if (tiles) {
        CompactKeptKernel<<<(unsigned)tiles, kKeepThreads, 0, s>>>(sorted, tuples, bits, part, canonical, tile_counts.ptr,
                                                                  col_indices.ptr, krow);
        GR_TRY(cudaGetLastError());
    }
    RowOffsetsKernel<SizeT><<<util::FillGrid(m + 1), 256, 0, s>>>(krow, m, local_nodes, row_offsets.ptr);
    GR_TRY(cudaGetLastError());
    GR_TRY(cudaStreamSynchronize(s));        // tile_counts / block_sums go out of scope
    if (kept_out) *kept_out = m;
    if (krow_out) *krow_out = krow;
    return cudaSuccess;
}

// bits needed to hold ids 0 .. nodes-1 (at least 1)
inline int IdBits(long long nodes)
{
    int b = 1;
    while ((1ll << b) < nodes) ++b;
    return b;
}

// keys of every entry of a resident CSR: (R << bits | C), R / C the (optionally mapped) row and
// column id, swapped when `transpose`.  A CTA takes 256 consecutive rows and walks their entries
// with all threads (a hub row is spread over the CTA, not walked by one thread).
template <typename SizeT>
__global__ void __launch_bounds__(256)
EdgeKeysKernel(const SizeT *row_offsets, const int *col, long long rows, const int *id_map, bool transpose, int bits,
               unsigned long long *keys)
{
    __shared__ long long s_off[257];
    for (long long r0 = (long long)blockIdx.x * 256; r0 < rows; r0 += (long long)gridDim.x * 256) {
        __syncthreads();
        const long long cnt = rows - r0 < 256 ? rows - r0 : 256;
        for (int t = threadIdx.x; t <= cnt; t += 256) s_off[t] = (long long)row_offsets[r0 + t];
        __syncthreads();
        const long long b = s_off[0], e = s_off[cnt];
        for (long long j = b + threadIdx.x; j < e; j += 256) {
            int lo = 0, hi = (int)cnt;               // last row with s_off[row] <= j
            while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (s_off[mid] <= j) lo = mid; else hi = mid; }
            long long R = r0 + lo, C = col[j];
            if (id_map) { R = id_map[R]; C = id_map[C]; }
            keys[j] = transpose ? (((unsigned long long)C << bits) | (unsigned long long)R)
                                : (((unsigned long long)R << bits) | (unsigned long long)C);
        }
    }
}

// the transpose (CSC) of a single-device CSR: csc_offsets / csc_indices of `g`.  Multiplicities
// and self loops are kept, so the pull direction inspects exactly the forward graph's entries.
template <typename SizeT>
cudaError_t BuildTranspose(DeviceCsr<int, SizeT> &g, cudaStream_t s)
{
    if (g.part.log_parts != 0)
        return util::GRError(cudaErrorInvalidValue, "transpose: single-device graphs only", __FILE__, __LINE__);
    if (g.streams_from_host)
        return util::GRError(cudaErrorInvalidValue, "transpose: the CSR is read in place from host memory", __FILE__, __LINE__);
    const unsigned long long m = (unsigned long long)g.edges;
    const int bits = IdBits(g.nodes);
    util::DeviceArray<unsigned long long> keys_a, keys_b;
    GR_TRY(keys_a.Allocate((size_t)m + 1));
    GR_TRY(keys_b.Allocate((size_t)m + 1));
    if (m) {
        EdgeKeysKernel<SizeT><<<util::FillGrid((unsigned long long)g.local_nodes), 256, 0, s>>>(
            g.row_offsets.ptr, g.col_indices.ptr, g.local_nodes, nullptr, true, bits, keys_a.ptr);
        GR_TRY(cudaGetLastError());
    }
    util::RadixSortScratch scratch;
    unsigned long long *sorted = keys_a.ptr;
    GR_TRY(util::RadixSort64(keys_a.ptr, keys_b.ptr, m, 2 * bits, scratch, &sorted, s));
    unsigned long long *spare = (sorted == keys_a.ptr) ? keys_b.ptr : keys_a.ptr;
    return SortedKeysToCsr<SizeT>(sorted, spare, m, bits, Partition(), false, g.local_nodes, g.csc_offsets,
                                  g.csc_indices, nullptr, nullptr, s);
}

// number of positions at which two offset arrays / two column arrays differ (symmetry check)
template <typename SizeT>
__global__ void CountDifferencesKernel(const SizeT *off_a, const SizeT *off_b, unsigned long long n_off,
                                       const int *col_a, const int *col_b, unsigned long long m,
                                       unsigned long long *diff)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long mine = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_off + m; i += stride) {
        if (i < n_off) mine += off_a[i] != off_b[i];
        else           mine += col_a[i - n_off] != col_b[i - n_off];
    }
    if (mine) atomicAdd(diff, mine);
}

// ---- degree order ------------------------------------------------------------------------------------
template <typename SizeT>
__global__ void DegreeKeysKernel(const SizeT *row_offsets, long long nodes, unsigned long long *keys)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nodes; v += stride) {
        unsigned long long d = (unsigned long long)(row_offsets[v + 1] - row_offsets[v]);
        if (d > 0xffffffffull) d = 0xffffffffull;
        keys[v] = ((0xffffffffull - d) << 32) | (unsigned long long)v;      // descending degree, ties by ascending id
    }
}

static __global__ void PermFromKeysKernel(const unsigned long long *sorted, long long nodes, int *to_new, int *to_old)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < nodes; r += stride) {
        const int old = (int)(sorted[r] & 0xffffffffull);
        to_old[r] = old;
        to_new[old] = (int)r;
    }
}

// out[v] = in[map[v]] (labels back to the caller's ids); MAP_VALUES: the values are vertex ids of the
// twin (predecessors) and are translated too, negative markers are kept
template <bool MAP_VALUES>
__global__ void GatherThroughKernel(const int *in, const int *to_new, const int *to_old, long long nodes, int *out)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nodes; v += stride) {
        int x = in[to_new[v]];
        if (MAP_VALUES && x >= 0) x = to_old[x];
        out[v] = x;
    }
}

// The degree-ordered twin of a single-device graph: the same graph with vertex ids renumbered by
// descending out-degree (ties by ascending id) and every row sorted by the new ids, i.e. hubs first.
// to_new[v] = id of v in the twin, to_old = its inverse.  A directed graph's twin gets its own transpose.
template <typename SizeT>
cudaError_t BuildDegreeOrdered(const DeviceCsr<int, SizeT> &g, DeviceCsr<int, SizeT> &twin,
                               util::DeviceArray<int> &to_new, util::DeviceArray<int> &to_old, cudaStream_t s)
{
    if (g.part.log_parts != 0 || g.streams_from_host)
        return util::GRError(cudaErrorInvalidValue, "degree order: resident single-device graphs only", __FILE__, __LINE__);
    const long long n = g.nodes;
    const unsigned long long m = (unsigned long long)g.edges;
    const int bits = IdBits(n);
    GR_TRY(to_new.Allocate((size_t)n + 1));
    GR_TRY(to_old.Allocate((size_t)n + 1));
    util::RadixSortScratch scratch;
    {
        util::DeviceArray<unsigned long long> ka, kb;
        GR_TRY(ka.Allocate((size_t)n + 1));
        GR_TRY(kb.Allocate((size_t)n + 1));
        unsigned long long *sorted = ka.ptr;
        if (n) {
            DegreeKeysKernel<SizeT><<<util::FillGrid((unsigned long long)n), 256, 0, s>>>(g.row_offsets.ptr, n, ka.ptr);
            GR_TRY(cudaGetLastError());
            GR_TRY(util::RadixSort64(ka.ptr, kb.ptr, (unsigned long long)n, 64, scratch, &sorted, s));
            PermFromKeysKernel<<<util::FillGrid((unsigned long long)n), 256, 0, s>>>(sorted, n, to_new.ptr, to_old.ptr);
            GR_TRY(cudaGetLastError());
        }
        GR_TRY(cudaStreamSynchronize(s));
    }
    util::DeviceArray<unsigned long long> keys_a, keys_b;
    GR_TRY(keys_a.Allocate((size_t)m + 1));
    GR_TRY(keys_b.Allocate((size_t)m + 1));
    if (m) {
        EdgeKeysKernel<SizeT><<<util::FillGrid((unsigned long long)g.local_nodes), 256, 0, s>>>(
            g.row_offsets.ptr, g.col_indices.ptr, g.local_nodes, to_new.ptr, false, bits, keys_a.ptr);
        GR_TRY(cudaGetLastError());
    }
    unsigned long long *sorted = keys_a.ptr;
    GR_TRY(util::RadixSort64(keys_a.ptr, keys_b.ptr, m, 2 * bits, scratch, &sorted, s));
    unsigned long long *spare = (sorted == keys_a.ptr) ? keys_b.ptr : keys_a.ptr;
    twin.nodes = twin.local_nodes = n; twin.part = Partition(); twin.device = g.device;
    twin.symmetric = g.symmetric; twin.edge_values.size = 0;
    unsigned long long kept = 0;
    GR_TRY(SortedKeysToCsr<SizeT>(sorted, spare, m, bits, Partition(), false, n, twin.row_offsets, twin.col_indices,
                                  &kept, nullptr, s));
    twin.edges = (long long)kept;
    keys_a.Release(); keys_b.Release();
    if (!g.symmetric && g.HasInverse()) GR_TRY(BuildTranspose(twin, s));
    return cudaSuccess;
}

}  // namespace graph
}  // namespace gunrock
