// stream_kernel.cuh -- the streaming instance of the filter operator for idempotent problems:
// big frontiers whose only cull is the exact visited-bitmask claim.
//
// Same operator as filter::Kernel (kernel.cuh; reference: gunrock/oprtr/filter/cta.cuh:166-306,
// 467-544 -- BitmaskCull with a texture fetch per item, VertexCull, history / warp hash, raking
// scan, one atomic per 512-item tile), rebuilt around what bounds it on a B200: the random probe of
// the visited map costs one L1 wavefront per item, ten times what streaming the item costs
// (tools/probe_micro.cu).  So, as in advance/lb_stream.cuh:
//   * the input frontier arrives in shared memory through TMA 1-D bulk copies (cp.async.bulk), one
//     16 KB tile per copy, kStreamStages tiles in flight per SM, one mbarrier per ring slot;
//   * the visited words of the hot vertices (ids < hot_words * 32; the high-degree ones on a
//     degree-ordered twin) are mirrored in shared memory and probed there; a clear bit sends the
//     item to the exact claim (atomicOr on the global word), whose winner is the unique survivor --
//     which makes the reference's label test, history hash and warp hash redundant;
//   * survivors are packed into a per-warp shared-memory stage with one ballot per four items and
//     leave with one global atomic per ~200 of them; there is no CTA-wide barrier inside the tile
//     loop (the generic kernel: one atomic per 1024-item tile behind two CTA barriers), warps meet
//     only at the ring's mbarriers and drift up to two tiles apart.
// One CTA of 1024 threads per SM.  The order of the output frontier is unspecified (the generic
// kernel keeps tiles stable; no caller depends on it).
//
// Restrictions (LaunchKernel of kernel.cuh is the general path): Problem::ENABLE_IDEMPOTENCE with
// the bitmask cull always on (END_BITMASK_CULL < 0), an explicit input queue, no predecessor queue,
// compaction, a single device.
#pragma once
#include "kernel.cuh"
#include "../../util/async_copy.cuh"
#include "../../graph/csr.cuh"

namespace gunrock {
namespace oprtr {
namespace filter {

constexpr int kStreamStages = 4;

constexpr int kStreamAhead = 2;         // tiles in flight ahead of the one being consumed
constexpr int kStreamWarpStage = 256;   // survivors a warp collects before it enqueues them (one global atomic)

template <typename Policy>
struct alignas(128) StreamStorage {
    int                ring[kStreamStages][Policy::TILE];
    int                wstage[Policy::THREADS / 32][kStreamWarpStage];
    unsigned long long full[kStreamStages];         // mbarriers: tile in slot s has landed
    unsigned long long empty[kStreamStages];        // mbarriers: every warp has left slot s
};
template <typename Policy>
constexpr size_t StreamHotOffset() { return (sizeof(StreamStorage<Policy>) + 127) & ~(size_t)127; }
template <typename Policy>
constexpr unsigned StreamMaxHotWords() { return (unsigned)((232448 - StreamHotOffset<Policy>()) / 4) & ~31u; }

template <typename Policy, typename Problem, typename Functor>
__global__ void __launch_bounds__(Policy::THREADS, 1)
StreamKernel(const __grid_constant__ typename Problem::DataSlice ds_in,
             const __grid_constant__ Args<typename Policy::VertexId, typename Problem::Value> args,
             const unsigned hot_words)
{
    typename Problem::DataSlice ds = ds_in;     // one device: compile-time identity partition (see lb_stream.cuh)
    ds.part = graph::Partition();
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef typename Problem::Value   Value;
    constexpr int THREADS = Policy::THREADS, TILE = Policy::TILE, N = kStreamStages;
    constexpr int QUADS = TILE / (4 * THREADS);
    static_assert(Problem::ENABLE_IDEMPOTENCE && Policy::END_BITMASK_CULL < 0, "bitmask-claim filter only");
    static_assert(TILE % (4 * THREADS) == 0 && sizeof(VertexId) == 4, "a thread consumes whole 16-byte groups");

    extern __shared__ __align__(128) unsigned char s_raw[];
    StreamStorage<Policy> &sm = *reinterpret_cast<StreamStorage<Policy> *>(s_raw);
    unsigned *s_hot = reinterpret_cast<unsigned *>(s_raw + StreamHotOffset<Policy>());
    const unsigned hot_bits = hot_words << 5;
    const int tid = threadIdx.x, lane = util::LaneId();
    const unsigned cta = blockIdx.x, nctas = gridDim.x;

    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (cta == 0 && tid == 0) ClearFutureSlot(args.slots, args.queue_index);
    if (tid == 0) {
        for (int s = 0; s < N; ++s) { util::MbarInit(&sm.full[s], 1); util::MbarInit(&sm.empty[s], THREADS / 32); }
        util::MbarInitFence();
    }
    for (unsigned w = tid; w < hot_words; w += THREADS) s_hot[w] = util::LoadCG(ds.d_visited_mask + w);
    __syncthreads();

    const unsigned long long num_tiles = (in_len + TILE - 1) / TILE;
    const long long my_tiles = num_tiles > cta ? (long long)((num_tiles - cta + nctas - 1) / nctas) : 0;
    // thread 0 is the producer: one bulk copy per tile (the queue is 16-byte aligned and padded)
    auto fill = [&](long long k) {
        const unsigned long long t = cta + (unsigned long long)k * nctas, base = t * TILE;
        const unsigned long long left = in_len - base;
        const unsigned items = left < (unsigned long long)TILE ? (unsigned)left : (unsigned)TILE;
        const unsigned bytes = ((items + 3u) & ~3u) * 4u;
        unsigned long long *bar = &sm.full[k % N];
        util::MbarExpectTx(bar, bytes);
        util::BulkCopyG2S(sm.ring[k % N], args.in_queue + base, bytes, bar);
        util::MbarArrive(bar);
    };
    int *wstage = sm.wstage[util::WarpId()];
    unsigned wcount = 0;                        // warp-uniform: entries in wstage
    auto flush_warp = [&]() {                   // warp-convergent
        if (wcount == 0) return;
        unsigned long long fb = 0;
        if (lane == 0) fb = atomicAdd(&slot_out->out_len, (unsigned long long)wcount);
        fb = __shfl_sync(util::kFullMask, fb, 0);
        __syncwarp();
        if (args.out_queue) {
            for (unsigned k = lane; k < wcount; k += 32) {
                if (fb + k < args.out_cap) args.out_queue[fb + k] = wstage[k];
                else slot_out->overflow = 1;
            }
        }
        __syncwarp();
        wcount = 0;
    };

    // no CTA-wide barrier inside the loop: thread 0 refills a slot once every warp has left it (`empty`),
    // consumers wait for their tile (`full`); warps drift up to N - AHEAD tiles apart
    constexpr int AHEAD = kStreamAhead;
    if (tid == 0)
        for (long long k = 0; k < AHEAD && k < my_tiles; ++k) fill(k);
    for (long long k = 0; k < my_tiles; ++k) {
        const long long ks = k + AHEAD;
        if (tid == 0 && ks < my_tiles) {
            if (ks >= N) util::MbarWait(&sm.empty[ks % N], (unsigned)(((ks / N) - 1) & 1));
            fill(ks);
        }
        __syncwarp();
        util::MbarWait(&sm.full[k % N], (unsigned)((k / N) & 1));
        const unsigned long long base = (cta + (unsigned long long)k * nctas) * TILE;
        const int4 *t4 = reinterpret_cast<const int4 *>(sm.ring[k % N]);
#pragma unroll
        for (int r = 0; r < QUADS; ++r) {
            const int qi = r * THREADS + tid;
            const int4 q = t4[qi];
            const VertexId v[4] = {q.x, q.y, q.z, q.w};
            bool live[4];
            unsigned word[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                live[i] = base + (unsigned long long)(qi * 4 + i) < in_len && v[i] >= 0;
                word[i] = 0;
                if (live[i])
                    word[i] = ((unsigned)v[i] < hot_bits) ? s_hot[(unsigned)v[i] >> 5]
                                                          : util::LoadCG(ds.d_visited_mask + (ds.BitPos(v[i]) >> 5));
            }
            // the four claims are independent: issue them together, then look at the answers
            unsigned old_word[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const unsigned bit = 1u << (ds.BitPos(v[i]) & 31);
                live[i] = live[i] && !(word[i] & bit);
                // the cached probe may be stale; the atomic's return value is exact: one claimant survives
                old_word[i] = util::AtomicOrIf(live[i], ds.d_visited_mask + (ds.BitPos(v[i]) >> 5), bit);
            }
            bool ok[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const unsigned bit = 1u << (ds.BitPos(v[i]) & 31);
                ok[i] = live[i] && !(old_word[i] & bit);
                if (live[i] && (unsigned)v[i] < hot_bits) atomicOr(s_hot + ((unsigned)v[i] >> 5), bit);
                if (ok[i]) {
                    const SizeT nid = (SizeT)(base + (unsigned long long)(qi * 4 + i));
                    if (Functor::CondFilter(v[i], &ds, (Value)args.iteration, nid))
                        Functor::ApplyFilter(v[i], &ds, (Value)args.iteration, nid);
                    else ok[i] = false;
                }
            }
            unsigned bal[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) bal[i] = __ballot_sync(util::kFullMask, ok[i]);
            const unsigned tot = __popc(bal[0]) + __popc(bal[1]) + __popc(bal[2]) + __popc(bal[3]);
            if (tot) {                                              // warp-uniform
                if (wcount + tot > (unsigned)kStreamWarpStage) flush_warp();
                const unsigned below = (1u << lane) - 1;
                unsigned at = wcount;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (ok[i]) wstage[at + __popc(bal[i] & below)] = v[i];
                    at += __popc(bal[i]);
                }
                wcount += tot;
            }
        }
        __syncwarp();
        if (lane == 0) util::MbarArrive(&sm.empty[k % N]);
    }
    flush_warp();
}

// callers check this before LaunchStreamKernel
template <typename Policy, typename Problem>
inline bool StreamEligible(const Args<typename Policy::VertexId, typename Problem::Value> &args, int log_parts)
{
    return Problem::ENABLE_IDEMPOTENCE && Policy::END_BITMASK_CULL < 0 && args.in_queue != nullptr &&
           args.in_pred_queue == nullptr && args.compact && log_parts == 0;
}

template <typename Policy, typename Problem, typename Functor>
cudaError_t LaunchStreamKernel(int sm_count, const typename Problem::DataSlice &ds,
                               const Args<typename Policy::VertexId, typename Problem::Value> &args,
                               unsigned hot_words, cudaStream_t stream)
{
    auto kernel = StreamKernel<Policy, Problem, Functor>;
    static bool configured = false;
    if (!configured) {
        GR_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
        configured = true;
    }
    if (hot_words > StreamMaxHotWords<Policy>()) hot_words = StreamMaxHotWords<Policy>();
    unsigned long long tiles = (args.in_len + Policy::TILE - 1) / Policy::TILE;
    unsigned grid = (unsigned)(tiles < 1 ? 1 : (tiles < (unsigned long long)sm_count ? tiles : (unsigned long long)sm_count));
    const size_t smem = StreamHotOffset<Policy>() + sizeof(unsigned) * (size_t)hot_words;
    kernel<<<grid, Policy::THREADS, smem, stream>>>(ds, args, hot_words);
    return util::GRError(cudaGetLastError(), "filter::StreamKernel launch failed", __FILE__, __LINE__);
}

}  // namespace filter
}  // namespace oprtr
}  // namespace gunrock
