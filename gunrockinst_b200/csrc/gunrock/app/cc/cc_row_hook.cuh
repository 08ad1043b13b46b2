// cc_row_hook.cuh -- the hook pass of connected components over rows binned by length.
//
// Same hook rule as HookFunctor (cc_functor.cuh; the reference's HookInit / HookMax,
// gunrock/app/cc/cc_functor.cuh:65-103,184-250: the larger root goes under the smaller one) and the
// same fixed point, so the labels after the flattening pass are identical.  What changes is who
// does the work.  Through the LB advance every CSR entry is a thread of its own: it pays the tile
// owner search, then chases BOTH end points to their roots (two dependent loads each once the
// giant component has formed).  Here a row stays with the lanes that walk it (oprtr/row_bins.cuh:
// four lanes, a warp, or a warp per 2048-entry piece), and every lane remembers the last root it
// found for the row's vertex: an ancestor stays an ancestor for ever (parents only move to
// ancestors, roots only get linked under smaller ids), so the next entry starts its chase from
// there, and an entry whose neighbour already hangs under that root (parent[d] == root: almost every
// entry once the giant component has formed) is done after ONE load, the neighbour's parent word.  The
// first hop of four neighbours is requested at once; chases read through L1 (see RootCached).
// GRB_CC_HOOK=lb selects the advance route.
#pragma once
#include "cc_functor.cuh"
#include "../../oprtr/row_bins.cuh"

namespace gunrock {
namespace app {
namespace cc {

namespace rows = oprtr::rows;

template <typename SizeT>
struct CcRowHookArgs {
    const SizeT *row_offsets;
    const int   *col_indices;
    int         *parent;
    bool         skip_upper;        // symmetric CSR: each undirected edge appears twice, hook it once
    rows::View   bins;
};

// The chases of this pass read parents through L1 (ld.global.ca).  A line in L1 can be stale -- another SM has
// moved a pointer meanwhile -- but every value a parent word ever held is an ancestor of its vertex (or the
// vertex itself, from the time it was a root), so a stale chase still walks upwards and ends at a vertex that
// was a root at some time; only the compare-and-swap (at L2) decides whether it still is.  What the cache buys:
// the root of the giant component is read by almost every hook, and at L2 that one sector serialises the pass
// (CcRowHookKernel with L2 loads took exactly as long as the LB advance: 12.8 of 14.4 ms at R-MAT scale 25).
__device__ __forceinline__ int RootCached(int *parent, int x, int p)     // p = a value of parent[x]
{
    while (p != x) {
        const int gp = util::LoadCA(parent + p);
        if (gp != p) util::StoreCG(parent + x, gp);     // path halving
        x = p; p = gp;
    }
    return x;
}

// link the components of u (a vertex that is or once was the root above the row's vertex) and of d, whose
// parent word read pd; returns the root both ended under.  No load at all when d already hangs under u.
__device__ __forceinline__ int HookPair(int *parent, int u, int d, int pd)
{
    if (pd == u) return u;                          // parent[d] == u: same tree, whatever u is by now
    int v = RootCached(parent, d, pd);
    if (v == u) return u;
    u = RootCached(parent, u, util::LoadCA(parent + u));
    for (;;) {
        if (u == v) return u;
        const int hi = u > v ? u : v, lo = u > v ? v : u;
        const int old = atomicCAS(parent + hi, hi, lo);
        if (old == hi) return lo;                   // linked larger root -> smaller root
        // `hi` stopped being a root meanwhile (old < hi): chase both sides again; max(u, v) strictly decreases
        u = RootCached(parent, old, util::LoadCG(parent + old));
        v = RootCached(parent, lo, util::LoadCG(parent + lo));
    }
}

// entries [beg, end) of row s by a whole warp; u = this lane's known ancestor of s
template <typename SizeT>
__device__ __forceinline__ void WarpHookRange(const CcRowHookArgs<SizeT> &a, int s, SizeT beg, SizeT end, int &u)
{
    const int *col = a.col_indices;
    SizeT e = beg + (SizeT)util::LaneId();
    for (; end - e > 96; e += 128) {
        int d[4], pd[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) d[j] = util::LoadStream(col + e + 32 * j);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const bool take = d[j] != s && !(a.skip_upper && d[j] > s);
            if (!take) d[j] = -1;
            pd[j] = take ? util::LoadCA(a.parent + d[j]) : 0;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (d[j] >= 0) u = HookPair(a.parent, u, d[j], pd[j]);
    }
    for (; e < end; e += 32) {
        const int d = util::LoadStream(col + e);
        if (d != s && !(a.skip_upper && d > s)) u = HookPair(a.parent, u, d, util::LoadCA(a.parent + d));
    }
}

template <typename SizeT>
__global__ void __launch_bounds__(rows::kThreads)
CcRowHookKernel(const __grid_constant__ CcRowHookArgs<SizeT> a)
{
    constexpr int WARPS = rows::kWarps;
    const rows::View &b = a.bins;
    const int lane = util::LaneId(), warp = util::WarpId();
    if (blockIdx.x < b.blocks_big) {
        const unsigned long long piece = (unsigned long long)blockIdx.x * WARPS + warp;
        if (piece < b.n_pieces) {
            SizeT beg, end;
            const int s = rows::PieceRange<SizeT>(b, a.row_offsets, piece, beg, end);
            int u = s;
            WarpHookRange<SizeT>(a, s, beg, end, u);
        }
    } else if (blockIdx.x < b.blocks_big + b.blocks_medium) {
        const unsigned long long first =
            ((unsigned long long)(blockIdx.x - b.blocks_big) * WARPS + warp) * (unsigned long long)rows::kMediumPerWarp;
        for (int k = 0; k < rows::kMediumPerWarp; ++k) {
            const unsigned long long idx = first + k;
            if (idx >= b.n_medium) break;                       // warp-uniform
            const int s = b.medium_rows[idx];
            int u = s;
            WarpHookRange<SizeT>(a, s, a.row_offsets[s], a.row_offsets[s + 1], u);
        }
    } else {
        const int group = lane >> 2, g = lane & 3;
        const unsigned long long first =
            ((unsigned long long)(blockIdx.x - b.blocks_big - b.blocks_medium) * WARPS + warp) *
            (unsigned long long)(8 * rows::kSmallIters);
        for (int k = 0; k < rows::kSmallIters; ++k) {
            const unsigned long long idx = first + (unsigned long long)(k * 8 + group);
            if (idx >= b.n_small) continue;
            const int s = b.small_rows[idx];
            const SizeT beg = a.row_offsets[s], end = a.row_offsets[s + 1];
            int d[rows::kSmallMax / 4], pd[rows::kSmallMax / 4];
#pragma unroll
            for (int j = 0; j < rows::kSmallMax / 4; ++j) {
                const SizeT e = beg + (SizeT)(g + 4 * j);
                d[j] = e < end ? util::LoadStream(a.col_indices + e) : -1;
            }
#pragma unroll
            for (int j = 0; j < rows::kSmallMax / 4; ++j) {
                const bool take = d[j] >= 0 && d[j] != s && !(a.skip_upper && d[j] > s);
                if (!take) d[j] = -1;
                pd[j] = take ? util::LoadCA(a.parent + d[j]) : 0;
            }
            int u = s;
#pragma unroll
            for (int j = 0; j < rows::kSmallMax / 4; ++j)
                if (d[j] >= 0) u = HookPair(a.parent, u, d[j], pd[j]);
        }
    }
}

// The same pass on a resident grid: warps draw a piece, 8 medium rows or 32 small rows per ticket (the next ticket is
// requested before the current unit is walked), a unit's row ids and offsets are loaded by one lane each.  The static
// grid above keeps a CTA's slots until its longest row is through (pr_row_sweep.cuh has the measurement).
constexpr int kHookMediumUnit = 8, kHookSmallUnit = 32;

template <typename SizeT>
__global__ void __launch_bounds__(rows::kThreads)
CcRowHookTicketKernel(const __grid_constant__ CcRowHookArgs<SizeT> a, unsigned long long *ticket)
{
    const rows::View &b = a.bins;
    const int lane = util::LaneId();
    const unsigned long long units_medium = (b.n_medium + kHookMediumUnit - 1) / kHookMediumUnit;
    const unsigned long long units_small = (b.n_small + kHookSmallUnit - 1) / kHookSmallUnit;
    const unsigned long long total = b.n_pieces + units_medium + units_small;
    unsigned long long t = 0;
    if (lane == 0) t = atomicAdd(ticket, 1ull);
    t = __shfl_sync(util::kFullMask, t, 0);
    while (t < total) {                                         // warp-uniform
        unsigned long long next = 0;
        if (lane == 0) next = atomicAdd(ticket, 1ull);
        if (t < b.n_pieces) {
            SizeT beg, end;
            const int s = rows::PieceRange<SizeT>(b, a.row_offsets, t, beg, end);
            int u = s;
            WarpHookRange<SizeT>(a, s, beg, end, u);
        } else if (t < b.n_pieces + units_medium) {
            const unsigned long long first = (t - b.n_pieces) * (unsigned long long)kHookMediumUnit;
            const int count = (int)(b.n_medium - first < (unsigned long long)kHookMediumUnit ? b.n_medium - first : kHookMediumUnit);
            int r = 0;
            SizeT beg = 0, end = 0;
            if (lane < count) { r = b.medium_rows[first + lane]; beg = a.row_offsets[r]; end = a.row_offsets[r + 1]; }
            for (int k = 0; k < count; ++k) {                   // warp-uniform
                const int s = __shfl_sync(util::kFullMask, r, k);
                const SizeT kb = __shfl_sync(util::kFullMask, beg, k), ke = __shfl_sync(util::kFullMask, end, k);
                int u = s;
                WarpHookRange<SizeT>(a, s, kb, ke, u);
            }
        } else {
            const unsigned long long first = (t - b.n_pieces - units_medium) * (unsigned long long)kHookSmallUnit;
            int r = 0;
            SizeT beg = 0, end = 0;
            if (first + (unsigned long long)lane < b.n_small) {
                r = b.small_rows[first + lane]; beg = a.row_offsets[r]; end = a.row_offsets[r + 1];
            }
            const int group = lane >> 2, g = lane & 3;
#pragma unroll 1
            for (int k = 0; k < 4; ++k) {
                const int s = __shfl_sync(util::kFullMask, r, 8 * k + group);
                const SizeT jb = __shfl_sync(util::kFullMask, beg, 8 * k + group);
                const SizeT je = __shfl_sync(util::kFullMask, end, 8 * k + group);
                int d[rows::kSmallMax / 4], pd[rows::kSmallMax / 4];
#pragma unroll
                for (int j = 0; j < rows::kSmallMax / 4; ++j) {
                    const SizeT e = jb + (SizeT)(g + 4 * j);
                    d[j] = e < je ? util::LoadStream(a.col_indices + e) : -1;
                }
#pragma unroll
                for (int j = 0; j < rows::kSmallMax / 4; ++j) {
                    const bool take = d[j] >= 0 && d[j] != s && !(a.skip_upper && d[j] > s);
                    if (!take) d[j] = -1;
                    pd[j] = take ? util::LoadCA(a.parent + d[j]) : 0;
                }
                int u = s;
#pragma unroll
                for (int j = 0; j < rows::kSmallMax / 4; ++j)
                    if (d[j] >= 0) u = HookPair(a.parent, u, d[j], pd[j]);
            }
        }
        t = __shfl_sync(util::kFullMask, next, 0);
    }
}

// The flattening pass of the row route: every vertex to its root, roots counted.  What PtrJumpFunctor does through
// the apply-only filter (the reference's PtrJump, gunrock/app/cc/cc_functor.cuh:116-170), with the hops read through
// L1 (roots do not move any more; a stale word still points upwards) and ONE atomic per CTA for the count instead
// of one per warp that holds a root (14 M components at R-MAT scale 25: about a million atomics on one word).
static __global__ void __launch_bounds__(256)
CcFlattenKernel(int *parent, long long nodes, unsigned long long *num_components)
{
    __shared__ unsigned long long s_red[256 / 32 + 1];
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned long long roots = 0;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nodes; v += stride) {
        int r = util::LoadCG(parent + v);
        if (r == (int)v) { ++roots; continue; }
        int p = util::LoadCA(parent + r);
        while (p != r) { r = p; p = util::LoadCA(parent + r); }
        util::StoreCG(parent + v, r);
    }
    oprtr::advance::BlockAccumulate<256>(roots, s_red, num_components);
}

}  // namespace cc
}  // namespace app
}  // namespace gunrock
