// kernel.cuh -- near/far priority queue: the Bisect operator.
//
// Counterpart of the reference's gunrock/priority_queue (kernel.cuh:116-246 MarkVisit /
// MarkNF / MarkValid / Compact / Compact2, host drivers Bisect :404-446 and RemoveInvalid
// :465-503, storage near_far_pile.cuh:38-140).  There a split is five launches, two
// moderngpu scans and two blocking 4-byte cudaMemcpys, and duplicates are resolved by a
// 4-byte-per-vertex "last writer wins" lookup table (d_visit_lookup).
//
// Here a split is ONE kernel: every item is scored with Functor::ComputePriorityScore
// (same contract as the reference's PQFunctor, app/sssp/sssp_functor.cuh:115-140) and goes
//     score <  lower            -> dropped (stale: it was already served by an earlier bucket)
//     lower <= score < upper    -> near queue (the next frontier)
//     upper <= score            -> far pile
// with both outputs compacted by one packed shuffle scan and two global atomics per
// 1024-item tile.  Duplicates in the far pile are prevented by a 1-bit-per-vertex
// membership map (test-and-set on entry, cleared on exit), so a pile never holds more than
// n ids (the reference sizes each pile at m + 1 ids, near_far_pile.cuh:96-125).
//
// Algorithmic HBM bytes per call: 4 (id) + 4 (priority word) per input item, 4 per item
// written.
#pragma once
#include "../oprtr/work_progress.cuh"
#include "../util/device_intrinsics.cuh"

namespace gunrock {
namespace priority_queue {

// storage of the far pile (ping-pong) and its membership bits
template <typename VertexId>
struct NearFarPile {
    util::DeviceArray<VertexId> far[2];
    util::DeviceArray<unsigned> far_mark;
    unsigned long long far_len = 0;
    int selector = 0;

    cudaError_t Allocate(size_t nodes)
    {
        GR_TRY(far[0].Allocate(nodes + 1));
        GR_TRY(far[1].Allocate(nodes + 1));
        return far_mark.Allocate((nodes + 31) / 32 + 1);
    }
    cudaError_t Reset(cudaStream_t s)
    {
        far_len = 0; selector = 0;
        return far_mark.Zero(s);
    }
};

template <typename VertexId>
struct BisectArgs {
    const VertexId     *in;             // items to split (-1 = hole)
    unsigned long long  in_len;         // host-known length, or an upper bound when d_in_len is set
    const unsigned long long *d_in_len; // optional device-side length
    VertexId           *near_out;       // next frontier
    unsigned long long  near_cap;
    VertexId           *far_out;        // far pile to append to
    unsigned long long  far_base;       // entries already in far_out
    unsigned long long  far_cap;
    unsigned           *far_mark;       // membership bits of the far pile
    bool                in_is_far;      // `in` IS the (other) far pile: leaving items clear their bit
    unsigned            lower, upper;   // bucket window [lower, upper)
    oprtr::LevelCounters *slots;
    long long           queue_index;    // counts land in slot queue_index + 1:
                                        //   out_len = near items, aux0 = far items appended,
                                        //   aux1 = ~(smallest score sent to the far pile)
};

template <int THREADS, int ITEMS, typename Problem, typename Functor>
__global__ void __launch_bounds__(THREADS)
BisectKernel(const __grid_constant__ typename Problem::DataSlice ds,
             const __grid_constant__ BisectArgs<typename Problem::VertexId> args)
{
    typedef typename Problem::VertexId VertexId;
    constexpr int TILE = THREADS * ITEMS;
    __shared__ unsigned long long s_warp[THREADS / 32 + 1];
    __shared__ unsigned long long s_near_base, s_far_base;

    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    oprtr::LevelCounters *slot = args.slots + ((args.queue_index + 1) & 3);
    if (blockIdx.x == 0 && threadIdx.x == 0) oprtr::ClearFutureSlot(args.slots, args.queue_index);

    unsigned min_far = 0xffffffffu;
    const unsigned long long tiles = (in_len + TILE - 1) / TILE;
    for (unsigned long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        VertexId id[ITEMS];
        unsigned char cls[ITEMS];           // 0 drop, 1 near, 2 far
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            const unsigned long long i = tile * TILE + (unsigned long long)r * THREADS + threadIdx.x;
            id[r] = (i < in_len) ? util::LoadStream(args.in + i) : (VertexId)-1;
        }
        unsigned long long packed = 0;      // near count in the low word, far count in the high word
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            cls[r] = 0;
            if (id[r] < 0) continue;
            const unsigned score = Functor::ComputePriorityScore(id[r], &ds);
            const unsigned long long pos = (unsigned long long)id[r];
            const unsigned bit = 1u << (pos & 31);
            if (score >= args.upper) {
                bool take = true;
                if (!args.in_is_far) take = (atomicOr(args.far_mark + (pos >> 5), bit) & bit) == 0;
                if (take) { cls[r] = 2; packed += 1ull << 32; min_far = score < min_far ? score : min_far; }
            } else {
                if (args.in_is_far) atomicAnd(args.far_mark + (pos >> 5), ~bit);
                if (score >= args.lower) { cls[r] = 1; packed += 1ull; }
            }
        }
        unsigned long long total;
        const unsigned long long excl = util::BlockExclusiveSum<THREADS>(packed, s_warp, total);
        if (total == 0) continue;                                  // uniform
        if (threadIdx.x == 0) {
            const unsigned long long tn = total & 0xffffffffull, tf = total >> 32;
            s_near_base = tn ? atomicAdd(&slot->out_len, tn) : 0ull;
            s_far_base  = tf ? atomicAdd(&slot->aux0, tf) : 0ull;
        }
        __syncthreads();
        unsigned long long at_near = s_near_base + (excl & 0xffffffffull);
        unsigned long long at_far  = args.far_base + s_far_base + (excl >> 32);
#pragma unroll
        for (int r = 0; r < ITEMS; ++r) {
            if (cls[r] == 1) {
                if (at_near < args.near_cap) args.near_out[at_near] = id[r]; else slot->overflow = 1;
                ++at_near;
            } else if (cls[r] == 2) {
                if (at_far < args.far_cap) args.far_out[at_far] = id[r]; else slot->overflow = 1;
                ++at_far;
            }
        }
        __syncthreads();                                            // bases are rewritten next tile
    }
    // smallest far score of the call: warp min, then one atomicMax of its complement per warp
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned t = __shfl_xor_sync(util::kFullMask, min_far, o);
        min_far = t < min_far ? t : min_far;
    }
    if (util::LaneId() == 0 && min_far != 0xffffffffu)
        atomicMax(&slot->aux1, (unsigned long long)(~min_far));
}

// host launcher (reference: priority_queue::Bisect, kernel.cuh:404-446)
template <typename Problem, typename Functor>
cudaError_t Bisect(int sm_count, const typename Problem::DataSlice &ds,
                   const BisectArgs<typename Problem::VertexId> &args, cudaStream_t stream)
{
    constexpr int THREADS = 256, ITEMS = 4;
    unsigned long long tiles = (args.in_len + THREADS * ITEMS - 1) / (THREADS * ITEMS);
    const unsigned long long resident = (unsigned long long)sm_count * 8;
    unsigned grid = (unsigned)(tiles < 1 ? 1 : (tiles < resident ? tiles : resident));
    BisectKernel<THREADS, ITEMS, Problem, Functor><<<grid, THREADS, 0, stream>>>(ds, args);
    return util::GRError(cudaGetLastError(), "priority_queue::BisectKernel launch failed", __FILE__, __LINE__);
}

}  // namespace priority_queue
}  // namespace gunrock
