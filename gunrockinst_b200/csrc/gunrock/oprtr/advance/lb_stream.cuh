// lb_stream.cuh -- the streaming instance of the load-balanced advance: big push levels of a
// search whose functor needs nothing but the neighbour id (BFS without predecessors).
//
// Replaces the same reference pipeline as lb.cuh (gunrock/oprtr/edge_map_partitioned/
// kernel.cuh:242-559, RelaxPartitionedEdges2: every thread fetches its neighbour id with a
// scalar (texture) load and probes / claims it in global memory).  What bounds that shape on a
// B200 is not DRAM but the L1 wavefront rate: one 128-byte line per lane for the random probe of
// the visited map (tools/probe_micro.cu: 334 G probes/s against 1650 G ids/s of plain
// streaming).  This kernel removes both per-edge global accesses from the SM's load/store path:
//
//  * neighbour-list tiles are STAGED IN SHARED MEMORY by the copy engines.  A tile is
//    TILE_EDGES consecutive entries of the scanned frontier's edge space (the split points come
//    from lb::ScanKernel, as for lb::ExpandKernel); it is made of pieces of neighbour lists.  A
//    piece of >= kStreamMinBulk entries arrives by ONE TMA 1-D bulk copy (cp.async.bulk, SASS
//    UBLKCP) issued by the thread that owns the piece -- the piece is rotated by up to three
//    slots inside its own slot range so that shared and global addresses agree modulo 16 bytes,
//    the <= 9 entries around the aligned run travel as 4-byte cp.async (LDGSTS); shorter pieces
//    are packed by their warp, coalesced, also with cp.async.  kStreamStages tiles are in flight
//    per SM; one mbarrier per ring slot collects the TMA transaction bytes and one deferred
//    arrival per thread (cp.async.mbarrier.arrive), so consumers wait on exactly one word.
//    Descriptors of the tile after next are prefetched into registers, so no thread ever waits
//    for a dependent global load between two tiles.
//  * the visited map of the HOT vertices -- ids below hot_words * 32 -- is copied into shared
//    memory when the CTA starts and probed there.  On the degree-ordered twin of a graph
//    (graph/build.cuh) the hot vertices are the highest-degree ones: 1 M bits (128 KB of the
//    228 KB an SM has) answer ~80 % of the probes of an R-MAT scale-25 search and ~97 % at scale
//    22.  A stale copy is harmless: a clear bit only sends the edge to the exact global claim.
//
// One CTA of Policy::THREADS (1024) threads per SM, persistent over its tiles, and no CTA-wide
// barrier inside the tile loop: warps meet only at the ring's mbarriers (`full`: a tile has landed,
// `empty`: every warp has left a slot), so they drift up to two tiles apart and one warp's L2
// round trips overlap the others' work.  Consumers read four ids per LDS.128, probe, issue their
// claims together (the functor's predicated CondEdgeIf) and enqueue through a per-warp
// shared-memory stage (one global atomic per ~200 discoveries).
//
// Restrictions (the launcher falls back to lb::ExpandKernel otherwise): V2V, OUT_COMPACT or
// OUT_NONE, !Problem::MARK_PREDECESSORS, a functor that ignores e_id / e_id_in, one device.
#pragma once
#include "lb.cuh"
#include "../../util/async_copy.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {
namespace lb {

#ifndef GRB_STREAM_COLD_LOAD
#define GRB_STREAM_COLD_LOAD util::LoadCG   // visited words outside the hot range: read at L2 (fresh: a claim another SM
#endif                                      // made in this launch is seen; a stale L1 line would cost an atomic per edge)
constexpr int kStreamStages  = 4;       // tiles in flight per SM
constexpr int kStreamMinBulk = 64;      // entries; shorter pieces are packed by their warp

constexpr int kStreamAhead   = 2;       // tiles staged ahead of the one being consumed; the other
                                        // kStreamStages - kStreamAhead slots let warps drift apart
constexpr int kStreamWarpStage = 256;   // discoveries a warp collects before it enqueues them (one global atomic)

template <typename Policy>
struct alignas(128) StreamStorage {
    int                   tile[kStreamStages][Policy::TILE_EDGES];
    int                   wstage[Policy::THREADS / 32][kStreamWarpStage];   // per-warp output staging
    unsigned long long    full[kStreamStages];          // mbarriers: tile in slot s has landed
    unsigned long long    empty[kStreamStages];         // mbarriers: every warp has left slot s
    unsigned long long    red[Policy::THREADS / 32 + 1];
    unsigned long long    total;
};
template <typename Policy>
constexpr size_t StreamHotOffset() { return (sizeof(StreamStorage<Policy>) + 127) & ~(size_t)127; }
// shared memory left for the hot part of the visited map, in 32-bit words
template <typename Policy>
constexpr unsigned StreamMaxHotWords() { return (unsigned)((232448 - StreamHotOffset<Policy>()) / 4) & ~31u; }

template <typename Policy, typename Problem, typename Functor>
__global__ void __launch_bounds__(Policy::THREADS, 1)
StreamExpandKernel(const __grid_constant__ typename Problem::DataSlice ds_in,
                   const __grid_constant__ Args<typename Policy::VertexId, typename Policy::SizeT> args,
                   const __grid_constant__ LbScratchView<typename Policy::SizeT> scratch,
                   const unsigned hot_words)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef unsigned long long        ScanT;
    constexpr int THREADS = Policy::THREADS, TILE = Policy::TILE_EDGES, N = kStreamStages;
    constexpr int QUADS = TILE / (4 * THREADS);
    static_assert(TILE % (4 * THREADS) == 0, "a thread consumes whole 16-byte groups");
    static_assert(!Problem::MARK_PREDECESSORS, "the streaming advance hands no source vertex to the functor");
    static_assert(sizeof(VertexId) == 4, "tiles hold 32-bit ids");

    // one device (the launcher checks): a compile-time identity partition drops the owner hash and the
    // owner-major bit positions from the per-edge code (as in pull::Sweep)
    typename Problem::DataSlice ds = ds_in;
    if constexpr (HasPartition<typename Problem::DataSlice>::value) ds.part = graph::Partition();

    extern __shared__ __align__(128) unsigned char s_raw[];
    StreamStorage<Policy> &sm = *reinterpret_cast<StreamStorage<Policy> *>(s_raw);
    unsigned *s_hot = reinterpret_cast<unsigned *>(s_raw + StreamHotOffset<Policy>());
    const unsigned hot_bits = hot_words << 5;
    const int tid = threadIdx.x, lane = util::LaneId();
    const unsigned cta = blockIdx.x, nctas = gridDim.x;

    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (cta == 0 && tid == 0) {
        ClearFutureSlot(args.slots, args.queue_index);
        *scratch.ticket = 0;                    // the scan of this call is complete (stream order)
    }
    if (tid == 0) {
        sm.total = scratch.scanned[in_len];
        for (int s = 0; s < N; ++s) { util::MbarInit(&sm.full[s], THREADS); util::MbarInit(&sm.empty[s], THREADS / 32); }
        util::MbarInitFence();
    }
    if constexpr (Policy::PRE_CULL) {
        for (unsigned w = tid; w < hot_words; w += THREADS) s_hot[w] = util::LoadCG(ds.d_visited_mask + w);
    }
    __syncthreads();
    const ScanT total = sm.total;
    unsigned long long num_tiles = (total + TILE - 1) / TILE;
    if (num_tiles > scratch.part_cap) {         // a frontier with duplicates expanded past the scratch
        num_tiles = scratch.part_cap;
        if (tid == 0) slot_out->overflow = 1;
    }
    const long long my_tiles = num_tiles > cta ? (long long)((num_tiles - cta + nctas - 1) / nctas) : 0;
    const VertexId *col = args.g.col_indices;
    unsigned long long my_edges = 0;            // thread 0 only
    unsigned long long my_degree = 0;           // TRACK_OUT_DEGREE: out-degrees of the vertices this thread enqueued
    int *wstage = sm.wstage[util::WarpId()];
    unsigned wcount = 0;                        // warp-uniform: entries in wstage
    // warp-convergent: the warp's staged discoveries leave for the output frontier as one coalesced run
    auto flush_warp = [&]() {
        if (wcount == 0) return;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&slot_out->out_len, (unsigned long long)wcount);
        base = __shfl_sync(util::kFullMask, base, 0);
        __syncwarp();
        for (unsigned k = lane; k < wcount; k += 32) {
            const VertexId u = wstage[k];
            if (base + k < args.out_cap) args.out_queue[base + k] = u;
            else slot_out->overflow = 1;
            if constexpr (Policy::TRACK_OUT_DEGREE) {
                const long long row = args.g.part.Local(u);
                my_degree += (unsigned long long)(args.g.row_offsets[row + 1] - args.g.row_offsets[row]);
            }
        }
        __syncwarp();
        wcount = 0;
    };

    // ---- staging ---------------------------------------------------------------------------------------
    // items [i0, i0 + cnt) overlap tile k of this CTA
    auto load_bounds = [&](long long k, long long &i0, long long &cnt) {
        const unsigned long long t = cta + (unsigned long long)k * nctas;
        i0 = scratch.part_start[t];
        const long long i1 = (t + 1 < num_tiles) ? scratch.part_start[t + 1] : (long long)in_len - 1;
        cnt = i1 - i0 + 1;
    };
    // item i0 + j: its range of the frontier's edge space and the CSR position of its first entry
    auto load_piece = [&](long long i0, long long cnt, long long j, ScanT &s, ScanT &e, SizeT &rs) {
        s = 0; e = 0; rs = 0;
        if (j < cnt) {
            s  = scratch.scanned[i0 + j];
            e  = scratch.scanned[i0 + j + 1];
            rs = scratch.row_start[i0 + j];
        }
    };
    // warp-convergent: copy this lane's piece (len entries from col + pos) to slots [d, d + len) of buf
    auto stage_piece = [&](int *buf, unsigned long long *bar, int d, SizeT pos, int len) {
        const bool is_long = len >= kStreamMinBulk;
        if (is_long) {
            // rotate the piece inside its slot range so that slot and CSR position agree modulo 4:
            // entries [0, main) go to slots d + shift ..., the last `shift` entries to slots d ...
            const int shift = (int)((pos - (SizeT)d) & 3);
            const int main_cnt = len - shift;
            const int head = (int)((4 - (int)(pos & 3)) & 3);       // entries before col's next 16-byte boundary
            const int bulk = (main_cnt - head) & ~3;
            int *dst = buf + d + shift;
            const VertexId *src = col + pos;
            for (int k = 0; k < head; ++k) util::CpAsync4(dst + k, src + k);
            if (bulk > 0) {
                util::MbarExpectTx(bar, (unsigned)bulk * 4u);
                util::BulkCopyG2S(dst + head, src + head, (unsigned)bulk * 4u, bar);
            }
            for (int k = head + bulk; k < main_cnt; ++k) util::CpAsync4(dst + k, src + k);
            for (int k = main_cnt; k < len; ++k) util::CpAsync4(buf + d + (k - main_cnt), src + k);
        }
        // short pieces of the warp: packed, so that consecutive lanes copy consecutive entries
        const int slen = is_long ? 0 : len;
        const int incl = util::WarpInclusiveSum(slen);
        const int T = __shfl_sync(util::kFullMask, incl, 31);
        const int pre = incl - slen;
        for (int x0 = 0; x0 < T; x0 += 32) {
            const int x = x0 + lane;
            int lo = 0;                                             // last lane whose exclusive prefix is <= x
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
                const int cand = lo + step;
                const int pc = __shfl_sync(util::kFullMask, pre, cand & 31);
                if (cand < 32 && pc <= x) lo = cand;
            }
            const int od = __shfl_sync(util::kFullMask, d, lo), opre = __shfl_sync(util::kFullMask, pre, lo);
            const long long opos = __shfl_sync(util::kFullMask, (long long)pos, lo);
            if (x < T) util::CpAsync4(buf + od + (x - opre), col + opos + (x - opre));
        }
    };
    // stage tile k into ring slot k % N; (s, e, rs) describe item i0 + tid (prefetched)
    auto issue = [&](long long k, long long i0, long long cnt, ScanT s, ScanT e, SizeT rs) {
        const unsigned long long t = cta + (unsigned long long)k * nctas;
        const ScanT tile_start = t * TILE;
        const ScanT tile_end = (total - tile_start < (ScanT)TILE) ? total : tile_start + TILE;
        int *buf = sm.tile[k % N];
        unsigned long long *bar = &sm.full[k % N];
        for (long long j0 = 0; j0 == 0 || j0 < cnt; j0 += THREADS) {       // one round unless > THREADS items overlap the tile
            if (j0 > 0) load_piece(i0, cnt, j0 + tid, s, e, rs);
            const ScanT b = s > tile_start ? s : tile_start, en = e < tile_end ? e : tile_end;
            const int len = (j0 + tid < cnt && en > b) ? (int)(en - b) : 0;
            stage_piece(buf, bar, (int)(b - tile_start), rs + (SizeT)(b - s), len);
        }
        util::MbarArriveOnCpAsync(bar);
        if (tid == 0) my_edges += tile_end - tile_start;
    };

    // ---- consumption -----------------------------------------------------------------------------------
    auto consume = [&](long long k) {
        const unsigned long long t = cta + (unsigned long long)k * nctas;
        const ScanT tile_start = t * TILE;
        const int tile_edges = (int)((total - tile_start < (ScanT)TILE) ? (total - tile_start) : (ScanT)TILE);
        const int4 *t4 = reinterpret_cast<const int4 *>(sm.tile[k % N]);
#pragma unroll
        for (int r = 0; r < QUADS; ++r) {
            const int qi = r * THREADS + tid;
            const int4 q = t4[qi];
            const VertexId v[4] = {q.x, q.y, q.z, q.w};
            bool live[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) live[i] = qi * 4 + i < tile_edges;
            if constexpr (Policy::PRE_CULL) {
                unsigned word[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    word[i] = 0;
                    if (live[i])
                        word[i] = ((unsigned)v[i] < hot_bits) ? s_hot[(unsigned)v[i] >> 5]
                                                              : GRB_STREAM_COLD_LOAD(ds.d_visited_mask + (ds.BitPos(v[i]) >> 5));
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if ((word[i] >> (ds.BitPos(v[i]) & 31)) & 1u) live[i] = false;
            }
            // all four conditions first: their claims (one global atomic each in BFS) are independent and
            // travel together; a thread that waited for each answer before asking the next question spent
            // four L2 round trips per tile on a level whose frontier discovers most of its neighbours
            bool ok[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                ok[i] = CondEdgeIfImpl<Functor, typename Problem::DataSlice, VertexId, SizeT>::Call(
                    live[i], (VertexId)args.iteration, v[i], &ds, (SizeT)0, (SizeT)0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (live[i]) {
                    // claimed now or found claimed: either way nobody in this CTA needs to ask again
                    if (Policy::PRE_CULL && (unsigned)v[i] < hot_bits) atomicOr(s_hot + ((unsigned)v[i] >> 5), 1u << (v[i] & 31));
                    if (ok[i]) Functor::ApplyEdge((VertexId)args.iteration, v[i], &ds, (SizeT)0, (SizeT)0);
                }
            }
            if (args.output == OUT_COMPACT) {
                unsigned bal[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) bal[i] = __ballot_sync(util::kFullMask, ok[i]);
                const unsigned tot = __popc(bal[0]) + __popc(bal[1]) + __popc(bal[2]) + __popc(bal[3]);
                if (tot) {                                          // warp-uniform
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
        }
    };

    // ---- the pipeline: tile k is consumed while tiles k+1 .. k+AHEAD are in flight.  No CTA-wide barrier inside:
    // a thread stages its share of tile k+AHEAD once every warp has left the slot's previous tenant (tile
    // k+AHEAD-N, `empty`), waits for its own tile on `full`, so warps drift up to N-AHEAD tiles apart and one
    // warp's L2 round trips (cold probes, claims) overlap the others' work. -----------------------------------
    constexpr int AHEAD = kStreamAhead;
    static_assert(AHEAD >= 1 && AHEAD < N, "ring depth");
    long long b_i0 = 0, b_cnt = 0, p_i0 = 0, p_cnt = 0;
    ScanT p_s = 0, p_e = 0;
    SizeT p_rs = 0;
    for (long long ks = 0; ks < AHEAD && ks < my_tiles; ++ks) {
        load_bounds(ks, p_i0, p_cnt);
        load_piece(p_i0, p_cnt, tid, p_s, p_e, p_rs);
        issue(ks, p_i0, p_cnt, p_s, p_e, p_rs);
    }
    if (AHEAD < my_tiles) { load_bounds(AHEAD, p_i0, p_cnt); load_piece(p_i0, p_cnt, tid, p_s, p_e, p_rs); }
    if (AHEAD + 1 < my_tiles) load_bounds(AHEAD + 1, b_i0, b_cnt);
    for (long long k = 0; k < my_tiles; ++k) {
        const long long ks = k + AHEAD;
        if (ks < my_tiles) {
            if (ks >= N) util::MbarWait(&sm.empty[ks % N], (unsigned)(((ks / N) - 1) & 1));
            issue(ks, p_i0, p_cnt, p_s, p_e, p_rs);
        }
        if (ks + 1 < my_tiles) { p_i0 = b_i0; p_cnt = b_cnt; load_piece(p_i0, p_cnt, tid, p_s, p_e, p_rs); }
        if (ks + 2 < my_tiles) load_bounds(ks + 2, b_i0, b_cnt);
        util::MbarWait(&sm.full[k % N], (unsigned)((k / N) & 1));
        consume(k);
        __syncwarp();
        if (lane == 0) util::MbarArrive(&sm.empty[k % N]);
    }
    if (args.output == OUT_COMPACT) flush_warp();
    BlockAccumulate<THREADS>(my_edges, sm.red, &slot_out->edges);
    if constexpr (Policy::TRACK_OUT_DEGREE) BlockAccumulate<THREADS>(my_degree, sm.red, &slot_out->out_degree);
}

}  // namespace lb
}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
