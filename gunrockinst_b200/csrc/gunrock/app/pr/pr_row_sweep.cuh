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
// walks never changes after the pruning phase, so the rows are binned ONCE by their CSR length
// (oprtr/row_bins.cuh: 1..16 entries, 17..2048, longer ones cut into 2048-entry pieces) and a sweep
// is a single launch over the bins.  The lane that ends up with a row's sum finishes the row on the
// spot (damping, moving test, new rank, new contribution into the other generation of the
// contribution vector), so there is no accumulator array, no atomic and no second pass for small
// and medium rows; only the big rows (a few thousand at scale 25) are finished by a follow-up
// kernel.  Sums of small and medium rows are taken in a fixed order (deterministic); big rows add
// their pieces atomically, as the generic route does.
//
// Two kernels: PrRowSweepKernel (static grid: CTAs by bin, the first version, GRB_PR_SWEEP=rows1;
// 7.4 ms per sweep) and PrRowSweepTicketKernel (resident grid, work drawn from a ticket counter,
// metadata and column ids requested ahead; the default, 4.9 ms, 4.0 ms on the degree-ordered twin).
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

// ---- the same sweep with tickets ------------------------------------------------------------------------
// The static grid above loses time at its end-of-CTA barrier (ncu, scale 25: 28 % of the warp samples sit there):
// the eight warps of a CTA get rows of very different lengths and the CTA's slots stay taken until the longest
// is through.  Here the grid is resident (CTAs per SM x SMs) and every warp draws its next unit of work from one
// counter -- pieces first, then units of 8 medium rows, then units of 32 small rows -- with the next ticket
// requested before the current unit is processed.  Inside a unit the rows' offsets, degrees and old ranks are
// loaded by one lane each up front (one coalesced request per array instead of a dependent chain per row), the
// first 128 column ids of the next medium row are requested before the current row's contributions are summed,
// all sixteen column ids a lane may need for its four small rows are requested at once, and rows are finished by
// the lane that loaded their metadata.
constexpr int kMediumUnit = 8;      // medium rows per ticket
constexpr int kSmallUnit  = 32;     // small rows per ticket

// up to four column ids of entries base + lane + 32 j (j < 4) of a row ending at `end`; -1 beyond the end
template <typename SizeT>
__device__ __forceinline__ void PrLoadCols(const int *col, SizeT base, SizeT end, int (&c)[4])
{
    const SizeT e = base + (SizeT)util::LaneId();
#pragma unroll
    for (int j = 0; j < 4; ++j) c[j] = (end - e > (SizeT)(32 * j)) ? util::LoadStream(col + e + 32 * j) : -1;
}

// sum over entries [beg, end) whose first round of column ids is already in c; result in every lane
template <typename SizeT>
__device__ __forceinline__ double PrGatherFrom(const PrRowSweepArgs<SizeT> &a, SizeT beg, SizeT end, const int (&c)[4])
{
    double s[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (c[j] >= 0) s[j] = __ldg(a.contrib_in + PrReplicaPos(a, c[j]));
    for (SizeT base = beg + 128; base < end; base += 128) {
        int d[4];
        PrLoadCols<SizeT>(a.col_indices, base, end, d);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (d[j] >= 0) s[j] += __ldg(a.contrib_in + PrReplicaPos(a, d[j]));
    }
    return util::WarpReduceSum((s[0] + s[1]) + (s[2] + s[3]));
}

// PrFinishRow with the row's degree and old rank already in registers
template <typename SizeT>
__device__ __forceinline__ bool PrFinishLoaded(const PrRowSweepArgs<SizeT> &a, int r, double sum, int d, double old_rank)
{
    const int v = (int)a.part.Global(r);
    const double base = (a.src == v || a.src == -1) ? 1.0 : 0.0;
    const double next = __dadd_rn(__dmul_rn(a.delta, sum), __dmul_rn(1.0 - a.delta, base));
    a.rank[r] = next;
    a.contrib_out[r] = next / (double)d;
    return fabs(next - old_rank) > a.threshold;
}

template <typename SizeT>
__global__ void __launch_bounds__(rows::kThreads)
PrRowSweepTicketKernel(const __grid_constant__ PrRowSweepArgs<SizeT> a, unsigned long long *ticket)
{
    __shared__ unsigned long long s_red[rows::kThreads / 32 + 1];
    const rows::View &b = a.bins;
    const int lane = util::LaneId();
    const unsigned long long units_medium = (b.n_medium + kMediumUnit - 1) / kMediumUnit;
    const unsigned long long units_small = (b.n_small + kSmallUnit - 1) / kSmallUnit;
    const unsigned long long total = b.n_pieces + units_medium + units_small;
    unsigned long long moving = 0;
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(ticket, 1ull);
    t = __shfl_sync(util::kFullMask, t, 0);
    while (t < total) {                                         // warp-uniform
        unsigned long long next = 0;
        if (lane == 0) next = atomicAdd(ticket, 1ull);          // answered while this unit is worked on
        if (t < b.n_pieces) {
            // ---- one 2048-entry piece of a big row
            SizeT beg, end;
            const int r = rows::PieceRange<SizeT>(b, a.row_offsets, t, beg, end);
            int c[4];
            PrLoadCols<SizeT>(a.col_indices, beg, end, c);
            const double sum = PrGatherFrom<SizeT>(a, beg, end, c);
            if (lane == 0 && sum != 0.0) atomicAdd(a.acc + r, sum);
        } else if (t < b.n_pieces + units_medium) {
            // ---- eight medium rows: lane k < 8 owns row k of the unit
            const unsigned long long first = (t - b.n_pieces) * (unsigned long long)kMediumUnit;
            const int count = (int)(b.n_medium - first < (unsigned long long)kMediumUnit ? b.n_medium - first : kMediumUnit);
            const bool own = lane < count;
            int r = 0, dg = 1;
            SizeT beg = 0, end = 0;
            double old_rank = 0.0, mine = 0.0;
            if (own) {
                r = b.medium_rows[first + lane];
                beg = a.row_offsets[r]; end = a.row_offsets[r + 1];
                dg = a.deg[r]; old_rank = a.rank[r];
            }
            SizeT kb = __shfl_sync(util::kFullMask, beg, 0), ke = __shfl_sync(util::kFullMask, end, 0);
            int c[4];
            PrLoadCols<SizeT>(a.col_indices, kb, ke, c);
            for (int k = 0; k < count; ++k) {
                int n[4] = {-1, -1, -1, -1};
                SizeT nb = 0, ne = 0;
                if (k + 1 < count) {                            // warp-uniform
                    nb = __shfl_sync(util::kFullMask, beg, k + 1); ne = __shfl_sync(util::kFullMask, end, k + 1);
                    PrLoadCols<SizeT>(a.col_indices, nb, ne, n);
                }
                const double sum = PrGatherFrom<SizeT>(a, kb, ke, c);
                if (lane == k) mine = sum;
                kb = nb; ke = ne;
#pragma unroll
                for (int j = 0; j < 4; ++j) c[j] = n[j];
            }
            if (own && PrFinishLoaded<SizeT>(a, r, mine, dg, old_rank)) ++moving;
        } else {
            // ---- 32 small rows: lane l owns row l of the unit; row 8 k + g is summed by lanes 4 g .. 4 g + 3 in round k
            const unsigned long long first = (t - b.n_pieces - units_medium) * (unsigned long long)kSmallUnit;
            const bool own = first + (unsigned long long)lane < b.n_small;
            int r = 0, dg = 1;
            SizeT beg = 0, end = 0;
            double old_rank = 0.0, mine = 0.0;
            if (own) {
                r = b.small_rows[first + lane];
                beg = a.row_offsets[r]; end = a.row_offsets[r + 1];
                dg = a.deg[r]; old_rank = a.rank[r];
            }
            const int group = lane >> 2, g = lane & 3;
            int d[4][rows::kSmallMax / 4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const SizeT jb = __shfl_sync(util::kFullMask, beg, 8 * k + group);
                const SizeT je = __shfl_sync(util::kFullMask, end, 8 * k + group);
#pragma unroll
                for (int j = 0; j < rows::kSmallMax / 4; ++j) {
                    const SizeT e = jb + (SizeT)(g + 4 * j);
                    d[k][j] = e < je ? util::LoadStream(a.col_indices + e) : -1;
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                double sum = 0.0;
#pragma unroll
                for (int j = 0; j < rows::kSmallMax / 4; ++j)
                    if (d[k][j] >= 0) sum += __ldg(a.contrib_in + PrReplicaPos(a, d[k][j]));
                sum += __shfl_xor_sync(util::kFullMask, sum, 1);
                sum += __shfl_xor_sync(util::kFullMask, sum, 2);
                const double got = __shfl_sync(util::kFullMask, sum, (lane & 7) * 4);
                if ((lane >> 3) == k) mine = got;
            }
            if (own && PrFinishLoaded<SizeT>(a, r, mine, dg, old_rank)) ++moving;
        }
        t = __shfl_sync(util::kFullMask, next, 0);
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
