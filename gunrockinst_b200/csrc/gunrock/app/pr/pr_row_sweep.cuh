// pr_row_sweep.cuh -- one sweep of the gather formulation of PageRank as ONE kernel over rows
// binned by length.
//
// The arithmetic is the reference's (gunrock/app/pr/pr_functor.cuh:32-100: the edge functor
// adds rank[s] / degree[s] into the destination, the vertex functor damps, adds the base term
// and tests |next - curr| > threshold) as mgpr_enactor.cuh restates it for a symmetric graph:
//     next[v] = delta * sum_{u in row(v)} contrib[u] + (1 - delta) * base(v),   contrib[u] = curr[u] / degree[u].
// The generic route (LB advance with MgPrGatherFunctor + MgPrFinishKernel) pays a degree scan, the
// owner search of the LB tiles, one fp64 atomic per row piece and a second pass over the rows:
// ~100 thread instructions per entry, 11 ms per sweep at R-MAT scale 25.  The set of rows a sweep
// walks never changes after the pruning phase, so here the rows are binned ONCE by their CSR length,
//     small   1 .. 16 entries        4 lanes per row (8 rows per warp),
//     medium  17 .. 2048             one warp per row, four column loads and four gathers in flight per lane,
//     big     > 2048                 cut into 2048-entry pieces, one warp per piece, one atomic per piece,
// and a sweep is a single launch whose CTAs take the big pieces first, then the medium and the small
// rows.  The lane that ends up with a row's sum finishes the row on the spot (damping, moving test,
// new rank, new contribution into the other generation of the contribution vector), so there is no
// accumulator array, no atomic and no second pass for small and medium rows; only the big rows
// (a few thousand at scale 25) are finished by a follow-up kernel.  Sums of small and medium rows are
// taken in a fixed order (deterministic); big rows add their pieces atomically, as the generic route does.
//
// Single device only (the partitioned enactor keeps the generic route: its finishing kernel is also
// the coalesced publication into the peers' replicas).  GRB_PR_SWEEP=lb selects the generic route.
#pragma once
#include "../../oprtr/row_bins.cuh"
#include "../../oprtr/work_progress.cuh"

namespace gunrock {
namespace app {
namespace pr {

namespace rows = oprtr::rows;

template <typename SizeT>
struct PrRowSweepArgs {
    const SizeT  *row_offsets;
    const int    *col_indices;
    const double *contrib_in;       // whole replica of the generation this sweep reads
    double       *contrib_out;      // this device's segment of the generation this sweep writes (indexed by local row)
    const int    *deg;              // reduced out-degrees (> 0 for every binned row)
    double       *rank;
    double       *acc;              // big rows only; zero on entry, zero again after PrBigRowFinishKernel
    rows::View    bins;             // the alive rows (oprtr/row_bins.cuh)
    double        delta, threshold;
    int           src;
    graph::Partition part;
    long long     local_nodes;
    oprtr::LevelCounters *counters;
};

template <typename SizeT>
__device__ __forceinline__ unsigned long long PrReplicaPos(const PrRowSweepArgs<SizeT> &a, int v)
{
    return (unsigned long long)a.part.Owner(v) * (unsigned long long)a.local_nodes + (unsigned long long)a.part.Local(v);
}

// sum of the contributions behind entries [beg, end) by a whole (converged) warp; result in every lane
template <typename SizeT>
__device__ __forceinline__ double PrWarpGather(const PrRowSweepArgs<SizeT> &a, SizeT beg, SizeT end)
{
    const int *col = a.col_indices;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    SizeT e = beg + (SizeT)util::LaneId();
    for (; end - e > 96; e += 128) {
        const int d0 = util::LoadStream(col + e);
        const int d1 = util::LoadStream(col + e + 32);
        const int d2 = util::LoadStream(col + e + 64);
        const int d3 = util::LoadStream(col + e + 96);
        s0 += __ldg(a.contrib_in + PrReplicaPos(a, d0));
        s1 += __ldg(a.contrib_in + PrReplicaPos(a, d1));
        s2 += __ldg(a.contrib_in + PrReplicaPos(a, d2));
        s3 += __ldg(a.contrib_in + PrReplicaPos(a, d3));
    }
    for (; e < end; e += 32) s0 += __ldg(a.contrib_in + PrReplicaPos(a, util::LoadStream(col + e)));
    return util::WarpReduceSum((s0 + s1) + (s2 + s3));
}

// the vertex functor of the power iteration (gunrock/app/pr/pr_functor.cuh:73-100) for local row r whose
// in-neighbours' contributions add up to `sum`; true while the row is still moving
template <typename SizeT>
__device__ __forceinline__ bool PrFinishRow(const PrRowSweepArgs<SizeT> &a, int r, double sum)
{
    const int d = a.deg[r];
    const int v = (int)a.part.Global(r);
    const double base = (a.src == v || a.src == -1) ? 1.0 : 0.0;
    const double next = __dadd_rn(__dmul_rn(a.delta, sum), __dmul_rn(1.0 - a.delta, base));
    const bool moving = fabs(next - a.rank[r]) > a.threshold;
    a.rank[r] = next;
    a.contrib_out[r] = next / (double)d;
    return moving;
}

template <typename SizeT>
__global__ void __launch_bounds__(rows::kThreads)
PrRowSweepKernel(const __grid_constant__ PrRowSweepArgs<SizeT> a)
{
    __shared__ unsigned long long s_red[rows::kThreads / 32 + 1];
    constexpr int WARPS = rows::kWarps;
    const rows::View &b = a.bins;
    const int lane = util::LaneId(), warp = util::WarpId();
    unsigned long long moving = 0;
    if (blockIdx.x < b.blocks_big) {
        // ---- one 2048-entry piece of a big row per warp
        const unsigned long long piece = (unsigned long long)blockIdx.x * WARPS + warp;
        if (piece < b.n_pieces) {
            SizeT beg, end;
            const int r = rows::PieceRange<SizeT>(b, a.row_offsets, piece, beg, end);
            const double sum = PrWarpGather<SizeT>(a, beg, end);
            if (lane == 0 && sum != 0.0) atomicAdd(a.acc + r, sum);
        }
    } else if (blockIdx.x < b.blocks_big + b.blocks_medium) {
        // ---- one medium row per warp and round
        const unsigned long long first =
            ((unsigned long long)(blockIdx.x - b.blocks_big) * WARPS + warp) * (unsigned long long)rows::kMediumPerWarp;
        for (int k = 0; k < rows::kMediumPerWarp; ++k) {
            const unsigned long long idx = first + k;
            if (idx >= b.n_medium) break;                       // warp-uniform
            const int r = b.medium_rows[idx];
            const double sum = PrWarpGather<SizeT>(a, a.row_offsets[r], a.row_offsets[r + 1]);
            if (lane == 0 && PrFinishRow<SizeT>(a, r, sum)) ++moving;
        }
    } else {
        // ---- eight small rows per warp and round, four lanes each
        const int group = lane >> 2, g = lane & 3;
        const unsigned long long first =
            ((unsigned long long)(blockIdx.x - b.blocks_big - b.blocks_medium) * WARPS + warp) *
            (unsigned long long)(8 * rows::kSmallIters);
        for (int k = 0; k < rows::kSmallIters; ++k) {
            const unsigned long long idx = first + (unsigned long long)(k * 8 + group);
            const bool valid = idx < b.n_small;
            int r = 0;
            SizeT beg = 0, end = 0;
            if (valid) { r = b.small_rows[idx]; beg = a.row_offsets[r]; end = a.row_offsets[r + 1]; }
            int d[rows::kSmallMax / 4];
#pragma unroll
            for (int j = 0; j < rows::kSmallMax / 4; ++j) {
                const SizeT e = beg + (SizeT)(g + 4 * j);
                d[j] = e < end ? util::LoadStream(a.col_indices + e) : -1;
            }
            double sum = 0.0;
#pragma unroll
            for (int j = 0; j < rows::kSmallMax / 4; ++j)
                if (d[j] >= 0) sum += __ldg(a.contrib_in + PrReplicaPos(a, d[j]));
            sum += __shfl_xor_sync(util::kFullMask, sum, 1);
            sum += __shfl_xor_sync(util::kFullMask, sum, 2);
            if (valid && g == 0 && PrFinishRow<SizeT>(a, r, sum)) ++moving;
        }
    }
    oprtr::advance::BlockAccumulate<rows::kThreads>(moving, s_red, &a.counters->aux0);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&a.counters->edges, b.entries);
}

// big rows: the pieces' sums are in acc
template <typename SizeT>
__global__ void __launch_bounds__(256)
PrBigRowFinishKernel(const __grid_constant__ PrRowSweepArgs<SizeT> a)
{
    __shared__ unsigned long long s_red[256 / 32 + 1];
    unsigned long long moving = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.bins.n_big; i += stride) {
        const int r = a.bins.big_rows[i];
        const double sum = a.acc[r];
        a.acc[r] = 0.0;
        if (PrFinishRow<SizeT>(a, r, sum)) ++moving;
    }
    oprtr::advance::BlockAccumulate<256>(moving, s_red, &a.counters->aux0);
}

}  // namespace pr
}  // namespace app
}  // namespace gunrock
