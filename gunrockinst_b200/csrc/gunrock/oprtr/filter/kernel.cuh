// kernel.cuh -- the filter (vertex map + compaction) operator.
//
// Counterpart of the reference's oprtr/filter (kernel.cuh:211-300,738-792 and
// cta.cuh:166-306,467-544): per input item
//     [idempotent problems, early iterations] BitmaskCull      cta.cuh:166-207
//     VertexCull  (label test or CondFilter/ApplyFilter)       cta.cuh:213-253
//     [idempotent problems] HistoryCull + WarpCull             cta.cuh:259-306
//     flag -> CTA scan -> one atomicAdd per tile -> scatter    cta.cuh:467-544
// and of its apply-only variant Kernel2 (kernel.cuh:302-383) used when
// `filtering_flag == false`.
//
// sm_100a rewrite: the raking smem scan is a shuffle scan, the per-warp hash
// cull is an exact `__match_any_sync` duplicate test, the visited mask is read
// as 32-bit words through L1 (a stale hit only costs a later exact test) and
// updated with a word atomicOr instead of the reference's texture fetch + racy
// byte read-modify-write, and a tile is 4 items per thread (1024 per CTA) with
// all loads of a tile issued before they are consumed.
//
// Functor contract (unchanged): CondFilter(node, DataSlice*, Value v, SizeT nid)
// decides whether `node` survives, ApplyFilter(...) runs on survivors.  `v` is
// the iteration number (or the predecessor when a predecessor queue is given
// and Problem::MARK_PREDECESSORS), `nid` the position in the input queue.
//
// Algorithmic HBM bytes per call: 4 per input item (+1 mask probe when
// culling by bitmask) + 4 per output item.
#pragma once
#ifndef GRB_FILTER_MASK_LOAD
#define GRB_FILTER_MASK_LOAD util::LoadCA
#endif
#include "../work_progress.cuh"
#include "../../util/device_intrinsics.cuh"

namespace gunrock {
namespace oprtr {
namespace filter {

template <
    typename _Problem,
    int _THREADS          = 256,
    int _ITEMS_PER_THREAD = 4,
    int _MIN_CTAS         = 4,
    int _END_BITMASK_CULL = 0,      // idempotent: bitmask-cull while iteration < this (<0: always; 0: never)
    int _HISTORY_SLOTS    = 2048>   // idempotent: CTA-wide history hash entries
struct KernelPolicy {
    typedef _Problem Problem;
    typedef typename Problem::VertexId VertexId;
    typedef typename Problem::SizeT    SizeT;
    static const int THREADS          = _THREADS;
    static const int ITEMS_PER_THREAD = _ITEMS_PER_THREAD;
    static const int TILE             = _THREADS * _ITEMS_PER_THREAD;
    static const int MIN_CTAS         = _MIN_CTAS;
    static const int END_BITMASK_CULL = _END_BITMASK_CULL;
    static const int HISTORY_SLOTS    = _HISTORY_SLOTS;
};

template <typename VertexId, typename Value>
struct Args {
    const VertexId     *in_queue;           // may contain -1 holes; NULL = the identity frontier 0..in_len-1
    const VertexId     *in_pred_queue;      // optional parallel queue (MARK_PREDECESSORS problems)
    VertexId           *out_queue;          // NULL: count survivors only
    unsigned long long  in_len;             // host-known length (or upper bound when d_in_len is set)
    const unsigned long long *d_in_len;     // optional device-side length
    unsigned long long  out_cap;
    LevelCounters      *slots;
    long long           queue_index;        // survivors are counted in slot queue_index + 1
    long long           iteration;
    bool                compact;            // false: apply-only map (reference Kernel2), out = in
};

template <typename Policy, typename Problem, typename Functor>
__global__ void __launch_bounds__(Policy::THREADS, Policy::MIN_CTAS)
Kernel(const __grid_constant__ typename Problem::DataSlice ds,
       const __grid_constant__ Args<typename Policy::VertexId, typename Problem::Value> args)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef typename Problem::Value   Value;
    constexpr int THREADS = Policy::THREADS, IPT = Policy::ITEMS_PER_THREAD, TILE = Policy::TILE;
    constexpr bool IDEMPOTENT = Problem::ENABLE_IDEMPOTENCE;

    constexpr int WARPS = THREADS / 32;
    __shared__ unsigned s_cnt[IPT * WARPS];
    __shared__ unsigned s_total;
    __shared__ unsigned long long s_base;
    __shared__ VertexId s_history[IDEMPOTENT ? Policy::HISTORY_SLOTS : 1];

    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (blockIdx.x == 0 && threadIdx.x == 0) ClearFutureSlot(args.slots, args.queue_index);

    if constexpr (IDEMPOTENT) {
        for (int h = threadIdx.x; h < Policy::HISTORY_SLOTS; h += THREADS) s_history[h] = -1;
        __syncthreads();
    }
    const int lane = util::LaneId();
    const unsigned long long num_tiles = (in_len + TILE - 1) / TILE;

    for (unsigned long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const unsigned long long base = tile * TILE;
        VertexId id[IPT];
        VertexId pred[IPT];
#pragma unroll
        for (int r = 0; r < IPT; ++r) {
            const unsigned long long i = base + (unsigned long long)r * THREADS + threadIdx.x;
            id[r] = (i < in_len) ? (args.in_queue ? util::LoadStream(args.in_queue + i) : (VertexId)i) : (VertexId)-1;
            pred[r] = (VertexId)args.iteration;
            if (Problem::MARK_PREDECESSORS && args.in_pred_queue && i < in_len)
                pred[r] = util::LoadStream(args.in_pred_queue + i);
        }
        // ---- apply-only map (reference Kernel2) --------------------------------------------
        if (!args.compact) {
#pragma unroll
            for (int r = 0; r < IPT; ++r) {
                const unsigned long long i = base + (unsigned long long)r * THREADS + threadIdx.x;
                if (id[r] >= 0) {
                    if (Functor::CondFilter(id[r], &ds, (Value)pred[r], (SizeT)i))
                        Functor::ApplyFilter(id[r], &ds, (Value)pred[r], (SizeT)i);
                }
                if (args.out_queue && i < in_len && i < args.out_cap) args.out_queue[i] = id[r];
            }
            continue;
        }
        // ---- culls ------------------------------------------------------------------------------
        if constexpr (IDEMPOTENT) {
            const bool use_mask = Policy::END_BITMASK_CULL < 0 || args.iteration < Policy::END_BITMASK_CULL;
            if (use_mask) {
                unsigned word[IPT];
#pragma unroll
                for (int r = 0; r < IPT; ++r)
                    word[r] = (id[r] >= 0) ? GRB_FILTER_MASK_LOAD(ds.d_visited_mask + (ds.BitPos(id[r]) >> 5)) : 0u;
#pragma unroll
                for (int r = 0; r < IPT; ++r) {
                    if (id[r] < 0) continue;
                    const unsigned bit = 1u << (ds.BitPos(id[r]) & 31);
                    // the cached probe may be stale; the atomic's return value is exact, so one
                    // claimant per vertex survives (the reference's byte RMW is best effort)
                    if ((word[r] & bit) || (atomicOr(ds.d_visited_mask + (ds.BitPos(id[r]) >> 5), bit) & bit))
                        id[r] = -1;
                }
            }
        }
#pragma unroll
        for (int r = 0; r < IPT; ++r) {
            if (id[r] < 0) continue;
            const unsigned long long i = base + (unsigned long long)r * THREADS + threadIdx.x;
            if constexpr (IDEMPOTENT) {
                // reference VertexCull, idempotent branch (filter/cta.cuh:218-238)
                if (util::LoadCG(ds.d_labels + ds.LocalRow(id[r])) != -1) id[r] = -1;
                else if (Functor::CondFilter(id[r], &ds, (Value)pred[r], (SizeT)i))
                    Functor::ApplyFilter(id[r], &ds, (Value)pred[r], (SizeT)i);
            } else {
                if (Functor::CondFilter(id[r], &ds, (Value)pred[r], (SizeT)i))
                    Functor::ApplyFilter(id[r], &ds, (Value)pred[r], (SizeT)i);
                else id[r] = -1;
            }
        }
        if constexpr (IDEMPOTENT) {
            // history hash (CTA-wide, lossy) then exact duplicate removal inside the warp
#pragma unroll
            for (int r = 0; r < IPT; ++r) {
                if (id[r] >= 0) {
                    const int h = (unsigned)id[r] % Policy::HISTORY_SLOTS;
                    if (s_history[h] == id[r]) id[r] = -1; else s_history[h] = id[r];
                }
                const unsigned peers = __match_any_sync(util::kFullMask, id[r]);
                if (id[r] >= 0 && (__ffs(peers) - 1) != lane) id[r] = -1;
            }
        }
        // ---- compaction: ballots -> scan of (round, warp) counts -> one atomic -> scatter -------
        // Stable: survivors keep their input order inside a tile (as the reference's raking scan
        // does, filter/cta.cuh:520-544) and each warp writes one contiguous run per round.
        unsigned bal[IPT];
#pragma unroll
        for (int r = 0; r < IPT; ++r) bal[r] = __ballot_sync(util::kFullMask, id[r] >= 0);
        if (lane == 0) {
#pragma unroll
            for (int r = 0; r < IPT; ++r) s_cnt[r * WARPS + util::WarpId()] = (unsigned)__popc(bal[r]);
        }
        __syncthreads();
        if (util::WarpId() == 0) {
            unsigned running = 0;
            for (int c = 0; c < IPT * WARPS; c += 32) {
                const unsigned v = (c + lane < IPT * WARPS) ? s_cnt[c + lane] : 0u;
                const unsigned incl = util::WarpInclusiveSum(v);
                if (c + lane < IPT * WARPS) s_cnt[c + lane] = running + incl - v;
                running += __shfl_sync(util::kFullMask, incl, 31);
            }
            if (lane == 0) {
                s_total = running;
                if (running) s_base = atomicAdd(&slot_out->out_len, (unsigned long long)running);
            }
        }
        __syncthreads();
        if (s_total != 0 && args.out_queue) {
            const unsigned long long ob = s_base;
#pragma unroll
            for (int r = 0; r < IPT; ++r) {
                if (id[r] >= 0) {
                    const unsigned long long at = ob + s_cnt[r * WARPS + util::WarpId()]
                                                + (unsigned)__popc(bal[r] & ((1u << lane) - 1));
                    if (at < args.out_cap) args.out_queue[at] = id[r];
                    else slot_out->overflow = 1;
                }
            }
        }
        __syncthreads();                                        // s_cnt / s_base are rewritten next tile
    }
}

template <typename Policy, typename Problem, typename Functor>
cudaError_t LaunchKernel(
    int sm_count,
    const typename Problem::DataSlice &ds,
    const Args<typename Policy::VertexId, typename Problem::Value> &args,
    cudaStream_t stream)
{
    const unsigned long long resident = (unsigned long long)sm_count * Policy::MIN_CTAS;
    unsigned long long tiles = (args.in_len + Policy::TILE - 1) / Policy::TILE;
    unsigned grid = (unsigned)(tiles < 1 ? 1 : (tiles < resident ? tiles : resident));
    Kernel<Policy, Problem, Functor><<<grid, Policy::THREADS, 0, stream>>>(ds, args);
    return util::GRError(cudaGetLastError(), "filter::Kernel launch failed", __FILE__, __LINE__);
}

}  // namespace filter
}  // namespace oprtr
}  // namespace gunrock
