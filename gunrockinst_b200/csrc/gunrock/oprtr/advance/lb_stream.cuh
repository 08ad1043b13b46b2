// lb_stream.cuh -- the streaming instance of the load-balanced advance: big push levels of a
// search whose functor needs nothing but the neighbour id (BFS without predecessors).
//
// Replaces the same reference pipeline as lb.cuh (gunrock/oprtr/edge_map_partitioned/
// kernel.cuh:242-559, RelaxPartitionedEdges2: every thread fetches its neighbour id with a
// scalar (texture) load and probes / claims it in global memory).  What bounds that shape on a
// B200 is not DRAM: it is the L1 wavefront rate of the random probe of the visited map (one
// 128-byte line per lane; tools/probe_micro.cu, profiles/r02b_probe_micro_s25.json: 334 G probes/s
// against 1660 G ids/s of plain streaming), ~100 instructions per edge of owner lookup and
// staging in lb::ExpandKernel, and the latency of three dependent round trips (neighbour id ->
// visited word -> claim) per edge with nothing to overlap them.  This kernel splits the work
// between producer and consumer warps of one 1024-thread CTA per SM:
//
//  * PRODUCER WARPS (8) turn the scanned frontier into DENSE TILES OF NEIGHBOUR IDS IN SHARED
//    MEMORY.  A tile is TILE_EDGES consecutive entries of the frontier's edge space (split points
//    from lb::ScanKernel, as for lb::ExpandKernel), made of pieces of neighbour lists; the
//    producers share a tile's items round-robin (lane <-> item) and only issue copies, they
//    never wait for data.  A piece of >= 32 entries arrives by ONE TMA 1-D bulk copy
//    (cp.async.bulk, SASS UBLKCP) issued by the lane that owns it -- the piece is rotated by up
//    to three slots inside its own slot range so that shared and global addresses agree modulo
//    16 bytes (a tile is a bag of ids: their order is irrelevant), the <= 9 entries around the
//    aligned run travel as 4-byte cp.async (LDGSTS).  Shorter pieces are flattened: a start-flag
//    bitmap of the batch (one word per lane, in registers) and its popcount prefix give every
//    slot its owner with two shuffles and a popc, and consecutive lanes copy consecutive entries
//    with cp.async.  One mbarrier per ring slot collects the TMA transaction bytes and one
//    deferred arrival per producer lane (cp.async.mbarrier.arrive).
//  * CONSUMER WARPS (24) wait on that one word, read eight ids per lane with two LDS.128, free
//    the slot, and probe: the visited map of the HOT vertices -- ids below hot_words * 32 -- lives
//    in shared memory (fetched by TMA bulk copies when the CTA starts), the other words are read
//    at L2, all eight under complementary predicates with no branch.  On the degree-ordered twin
//    of a graph (graph/build.cuh) the hot vertices are the highest-degree ones and answer most
//    probes of an R-MAT search without touching the L1 / L2 path.  A stale copy is harmless: a
//    clear bit only makes the vertex a candidate.
//  * CLAIMS ARE REDUCTIONS, WINNERS ARE HARVESTED.  The claim of BFS is the flip of the vertex's bit in
//    Problem's visited mask (BFSFunctor::CondEdge: atomicOr, the caller whose call changed the bit
//    wins).  Waiting for that answer -- and then storing the label, staging the vertex and looking up
//    its degree -- is a chain of dependent L2 round trips that cost two thirds of a level
//    (3.2 ms against 1.3 ms without it at scale 25).  Here a neighbour that
//    looks unvisited gets its bit set by a fire-and-forget reduction (RED.OR on the global word, and
//    on the shared copy for a hot vertex, so that an SM asks once), nothing in the streaming loop
//    depends on memory beyond the probe, and the WINNERS of the level are read off afterwards by
//    HarvestKernel: the bits that are set now and were not before the level (a snapshot of the mask
//    taken by the launcher) -- each exactly once, in vertex order, so Functor::ApplyEdge (the label
//    store), the output frontier and the out-degree sum become coalesced passes over 1 bit per
//    vertex instead of random accesses per discovery.  The output frontier comes out sorted.
//
// Restrictions (the launcher falls back to lb::ExpandKernel otherwise): V2V, OUT_COMPACT or
// OUT_NONE, !Problem::MARK_PREDECESSORS, !Problem::ENABLE_IDEMPOTENCE (the functor's claim must be
// the visited bit), a functor that ignores s_id / e_id / e_id_in in ApplyEdge, one device.
#pragma once
#include "lb.cuh"
#include "../../util/async_copy.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {
namespace lb {

constexpr int kStreamProducers = 8;         // producer warps; the other warps of the CTA consume
constexpr int kStreamStages    = 3;         // tiles in the ring
constexpr int kStreamRefresh   = 4;         // tiles between two fetches of the hot words
constexpr int kStreamLong      = 32;        // pieces of at least this many entries travel by TMA bulk copy
constexpr unsigned kStreamHotCap = 49152;   // words: 1.5 M bits; beyond that the L1 that is left starves the cold probes

template <typename Policy>
struct alignas(128) StreamStorage {
    static const int CONSUMERS = Policy::THREADS / 32 - kStreamProducers;
    int                   tile[kStreamStages][Policy::TILE_EDGES];          // the ring
    long long             base[kStreamProducers][32];                       // batch of short pieces: CSR position of slot 0 of
    int                   dbase[kStreamProducers][32];                      //   ranked piece r minus its prefix; same for its tile slot
    unsigned              flags[kStreamProducers][32];                      //   start flags
    unsigned long long    full[kStreamStages];                              // mbarriers: tile in slot s has landed
    unsigned long long    empty[kStreamStages];                             // mbarriers: every consumer has read slot s
    unsigned long long    red[Policy::THREADS / 32 + 1];
    unsigned long long    total;
    unsigned long long    bar;                                              // mbarrier of the hot-map bulk copy
};
template <typename Policy>
constexpr size_t StreamHotOffset() { return (sizeof(StreamStorage<Policy>) + 127) & ~(size_t)127; }
// shared memory left for the hot part of the visited map, in 32-bit words
template <typename Policy>
constexpr unsigned StreamMaxHotWords()
{
    const unsigned fit = (unsigned)((232448 - StreamHotOffset<Policy>()) / 4) & ~31u;
    return fit < kStreamHotCap ? fit : kStreamHotCap;
}

// the visited word of a vertex: from the shared copy when `hot`, else from L2 when `cold` (fresh: a bit another SM set in
// this launch is seen, which spares a reduction); ~0 ("visited") when neither.  Two loads under complementary predicates
// into one register: no branch, so a lane's eight probes are in flight together.
__device__ __forceinline__ unsigned ProbeVisitedWord(bool hot, bool cold, unsigned shared_addr, const unsigned *global_addr)
{
    unsigned w = 0xffffffffu;
    asm volatile("{\n .reg .pred ph, pc;\n setp.ne.u32 ph, %3, 0;\n setp.ne.u32 pc, %4, 0;\n"
                 " @ph ld.shared.u32 %0, [%1];\n @pc ld.global.cg.u32 %0, [%2];\n}\n"
                 : "+r"(w) : "r"(shared_addr), "l"(global_addr), "r"((unsigned)hot), "r"((unsigned)cold) : "memory");
    return w;
}

// DIRECT = true is the streaming instance of the FILTER operator (filter::LaunchStreamKernel): the "edge space" is the
// input queue itself -- one TMA bulk copy per tile, no scan, no pieces -- holes (-1) are skipped, and the harvest applies
// Functor::CondFilter / ApplyFilter to the winners instead of ApplyEdge.  Everything else (ring, probes, claims) is shared.
template <typename Policy, typename Problem, typename Functor, bool DIRECT = false>
__global__ void __launch_bounds__(Policy::THREADS, 1)
StreamExpandKernel(const __grid_constant__ typename Problem::DataSlice ds_in,
                   const __grid_constant__ Args<typename Policy::VertexId, typename Policy::SizeT> args,
                   const __grid_constant__ LbScratchView<typename Policy::SizeT> scratch,
                   const unsigned hot_words)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef unsigned long long        ScanT;
    constexpr int THREADS = Policy::THREADS, WARPS = THREADS / 32, TILE = Policy::TILE_EDGES, N = kStreamStages;
    constexpr int PW = kStreamProducers, CW = WARPS - PW;
    static_assert(TILE == CW * 32 * 8, "a consumer lane takes eight ids of every tile");
    static_assert(!Problem::MARK_PREDECESSORS, "the streaming advance hands no source vertex to the functor");
    static_assert(Policy::PRE_CULL, "the streaming advance culls against Problem's visited bitmask");
    static_assert(sizeof(VertexId) == 4, "tiles hold 32-bit ids");

    // one device (the launcher checks): a compile-time identity partition drops the owner hash and the
    // owner-major bit positions from the per-edge code (as in pull::Sweep): bit position of vertex v is v
    typename Problem::DataSlice ds = ds_in;
    if constexpr (HasPartition<typename Problem::DataSlice>::value) ds.part = graph::Partition();

    extern __shared__ __align__(128) unsigned char s_raw[];
    StreamStorage<Policy> &sm = *reinterpret_cast<StreamStorage<Policy> *>(s_raw);
    unsigned *s_hot = reinterpret_cast<unsigned *>(s_raw + StreamHotOffset<Policy>());
    const unsigned s_hot_addr = util::SharedAddr(s_hot);
    const unsigned hot_bits = hot_words << 5;
    const int tid = threadIdx.x, lane = util::LaneId(), warp = util::WarpId();
    const unsigned cta = blockIdx.x, nctas = gridDim.x;
    const unsigned below = (1u << lane) - 1;

    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (cta == 0 && tid == 0) {
        ClearFutureSlot(args.slots, args.queue_index);
        if (!DIRECT) *scratch.ticket = 0;       // the scan of this call is complete (stream order)
    }
    if (tid == 0) {
        sm.total = DIRECT ? in_len : scratch.scanned[in_len];
        util::MbarInit(&sm.bar, 1);
        for (int s = 0; s < N; ++s) { util::MbarInit(&sm.full[s], DIRECT ? PW * 32 : PW * 32 + 1); util::MbarInit(&sm.empty[s], CW); }
        util::MbarInitFence();
    }
    __syncthreads();
    if (hot_words) {
        if (tid == 0) {                         // the hot words of the visited map: TMA 1-D bulk copies, 32 KB each
            // (whole 16-byte groups: the map's allocation is padded, bits past the last vertex are never probed)
            const unsigned copy_words = (hot_words + 3u) & ~3u;
            util::MbarExpectTx(&sm.bar, copy_words * 4u);
            for (unsigned w = 0; w < copy_words; w += 8192) {
                const unsigned cnt = copy_words - w < 8192u ? copy_words - w : 8192u;
                util::BulkCopyG2S(s_hot + w, ds.d_visited_mask + w, cnt * 4u, &sm.bar);
            }
            util::MbarArrive(&sm.bar);
        }
        if (warp >= PW) util::MbarWait(&sm.bar, 0);             // only the consumers read it
    }
    const ScanT total = sm.total;
    unsigned long long num_tiles = (total + TILE - 1) / TILE;
    if (!DIRECT && num_tiles > scratch.part_cap) {      // a frontier with duplicates expanded past the scratch
        num_tiles = scratch.part_cap;
        if (tid == 0) slot_out->overflow = 1;
    }
    const long long my_tiles = num_tiles > cta ? (long long)((num_tiles - cta + nctas - 1) / nctas) : 0;
    const VertexId *col = args.g.col_indices;
    unsigned long long my_edges = 0;            // one consumer thread

    if (warp < PW) {
      if constexpr (DIRECT) {
        // ================================ producers, filter instance ========================================
        // the tiles are consecutive stretches of the input queue: one bulk copy each (the queue's allocation is 16-byte
        // aligned and padded), issued by the first producer thread; the other producer lanes only keep the barrier's count
        unsigned refresh_phase = 0;
        for (long long k = 0; k < my_tiles; ++k) {
            const unsigned long long t = cta + (unsigned long long)k * nctas;
            const ScanT tile_start = t * TILE;
            const unsigned items = (unsigned)((total - tile_start < (ScanT)TILE) ? (total - tile_start) : (ScanT)TILE);
            unsigned long long *bar = &sm.full[k % N];
            if (hot_words && tid == 0 && k > 0 && (k % kStreamRefresh) == 0 && util::MbarTryWait(&sm.bar, refresh_phase & 1u)) {
                ++refresh_phase;
                const unsigned copy_words = (hot_words + 3u) & ~3u;
                util::MbarExpectTx(&sm.bar, copy_words * 4u);
                for (unsigned w = 0; w < copy_words; w += 8192) {
                    const unsigned cnt = copy_words - w < 8192u ? copy_words - w : 8192u;
                    util::BulkCopyG2S(s_hot + w, ds.d_visited_mask + w, cnt * 4u, &sm.bar);
                }
                util::MbarArrive(&sm.bar);
            }
            if (k >= N) util::MbarWait(&sm.empty[k % N], (unsigned)(((k / N) - 1) & 1));
            if (tid == 0) {
                const unsigned bytes = ((items + 3u) & ~3u) * 4u;
                util::MbarArriveExpectTx(bar, bytes);           // arrival and announcement in one: the copy cannot be missed
                util::BulkCopyG2S(sm.tile[k % N], args.in_queue + tile_start, bytes, bar);
            } else {
                util::MbarArrive(bar);
            }
        }
      } else {
        // ================================ producers ==========================================================
        // tile k of this CTA is global tile cta + k * nctas and lives in ring slot k % N; all producers work on the same
        // tile, its items dealt round-robin to their lanes.  Nothing here waits for data.
        auto tile_items = [&](long long k, long long &i0, long long &i1) {
            i0 = 0; i1 = -1;
            if (k < my_tiles) {
                const unsigned long long t = cta + (unsigned long long)k * nctas;
                i0 = scratch.part_start[t];
                i1 = (t + 1 < num_tiles) ? scratch.part_start[t + 1] : (long long)in_len - 1;     // inclusive
            }
        };
        auto item_desc = [&](long long i, long long i1, ScanT &s, ScanT &e, SizeT &rs) {
            s = 0; e = 0; rs = 0;
            if (i <= i1) { s = scratch.scanned[i]; e = scratch.scanned[i + 1]; rs = scratch.row_start[i]; }
        };
        long long i0, i1, n_i0, n_i1;
        unsigned refresh_phase = 0;             // thread 0: completed fetches of the hot words (phase 0 is the first one)
        tile_items(0, i0, i1);
        for (long long k = 0; k < my_tiles; ++k) {
            const unsigned long long t = cta + (unsigned long long)k * nctas;
            const ScanT tile_start = t * TILE;
            const ScanT tile_end = (total - tile_start < (ScanT)TILE) ? total : tile_start + TILE;
            int *buf = sm.tile[k % N];
            unsigned long long *bar = &sm.full[k % N];
            tile_items(k + 1, n_i0, n_i1);                      // requested now, needed for the next tile
            ScanT s, e, ns, ne;
            SizeT rs, nrs;
            item_desc(i0 + warp + PW * lane, i1, s, e, rs);
            // every kStreamRefresh tiles the shared copy of the hot words is fetched again (nobody waits for it: a word that
            // is overwritten with an older value only costs a reduction): what the other SMs have discovered meanwhile stops
            // being "unvisited" here, which is most of what a hot vertex costs -- without it every SM flips every hot vertex
            // itself, ~150 reductions per vertex instead of a few
            if (hot_words && tid == 0 && k > 0 && (k % kStreamRefresh) == 0 && util::MbarTryWait(&sm.bar, refresh_phase & 1u)) {
                ++refresh_phase;
                const unsigned copy_words = (hot_words + 3u) & ~3u;
                util::MbarExpectTx(&sm.bar, copy_words * 4u);
                for (unsigned w = 0; w < copy_words; w += 8192) {
                    const unsigned cnt = copy_words - w < 8192u ? copy_words - w : 8192u;
                    util::BulkCopyG2S(s_hot + w, ds.d_visited_mask + w, cnt * 4u, &sm.bar);
                }
                util::MbarArrive(&sm.bar);
            }
            // every consumer has read the slot's previous tenant (which also means that tenant's phase of `full` is over)
            if (k >= N) util::MbarWait(&sm.empty[k % N], (unsigned)(((k / N) - 1) & 1));
            for (long long ib = i0; ib <= i1; ib += 32 * PW) {
                // items are dealt round-robin to the producers, so that a tile of a few dozen long lists keeps all of them busy
                // (a warp issues its lanes' bulk copies one after the other)
                const long long i = ib + warp + PW * lane;
                item_desc(i + 32 * PW, i1, ns, ne, nrs);        // this producer's next batch, requested before this one is issued
                // lane <-> item: its range of the frontier's edge space clipped to the tile, where that starts in the CSR
                // and in the tile
                const ScanT b = s > tile_start ? s : tile_start, en = e < tile_end ? e : tile_end;
                const int len = (i <= i1 && en > b) ? (int)(en - b) : 0;
                const long long pos = (long long)rs + (long long)(b - s);
                const int d = (int)(b - tile_start);
                if (len >= kStreamLong) {
                    // rotate the piece inside its slot range so that slot and CSR position agree modulo 4:
                    // entries [0, main) go to slots d + shift ..., the last `shift` entries to slots d ...
                    const int shift = (int)((pos - (long long)d) & 3);
                    const int main_cnt = len - shift;
                    const int head = (int)((4 - (int)(pos & 3)) & 3);       // entries before col's next 16-byte boundary
                    const int bulk = (main_cnt - head) & ~3;
                    int *dst = buf + d + shift;
                    const VertexId *src = col + pos;
                    util::MbarExpectTx(bar, (unsigned)bulk * 4u);           // len >= 32: bulk > 0
                    util::BulkCopyG2S(dst + head, src + head, (unsigned)bulk * 4u, bar);
                    // the <= 3 entries before the aligned run and the <= 6 behind it (no loops: lanes differ)
#pragma unroll
                    for (int x = 0; x < 3; ++x)
                        if (x < head) util::CpAsync4(dst + x, src + x);
#pragma unroll
                    for (int y = 0; y < 6; ++y) {
                        const int x = head + bulk + y;
                        if (x < len) util::CpAsync4(x < main_cnt ? dst + x : buf + d + (x - main_cnt), src + x);
                    }
                }
                __syncwarp();
                // short pieces, flattened: slot x of the batch belongs to the piece whose start flag is the last one at or
                // before x; flag word w sits in lane w's register, so is the number of flags before it
                const int slen = len < kStreamLong ? len : 0;
                const unsigned nz = __ballot_sync(util::kFullMask, slen > 0);
                if (nz) {
                    const int incl = util::WarpInclusiveSum(slen);
                    const int T = __shfl_sync(util::kFullMask, incl, 31);   // <= 31 * 32: 31 flag words
                    const int pre = incl - slen;
                    sm.flags[warp][lane] = 0;
                    __syncwarp();
                    if (slen > 0) {
                        atomicOr(&sm.flags[warp][pre >> 5], 1u << (pre & 31));
                        const int r = __popc(nz & below);
                        sm.base[warp][r] = pos - pre;
                        sm.dbase[warp][r] = d - pre;
                    }
                    __syncwarp();
                    const unsigned fw = sm.flags[warp][lane];
                    const int fc = __popc(fw);
                    const int fpre = util::WarpInclusiveSum(fc) - fc;
                    for (int x0 = 0; x0 < T; x0 += 32) {
                        const int w = x0 >> 5;                              // warp-uniform
                        const unsigned word = __shfl_sync(util::kFullMask, fw, w);
                        const int wp = __shfl_sync(util::kFullMask, fpre, w);
                        const int x = x0 + lane;
                        if (x < T) {
                            const int rank = wp + __popc(word & (0xffffffffu >> (31 - lane))) - 1;
                            util::CpAsync4(buf + (sm.dbase[warp][rank] + x), col + (sm.base[warp][rank] + x));
                        }
                    }
                    __syncwarp();                                           // flags / base are rewritten by the next batch
                }
                s = ns; e = ne; rs = nrs;
            }
            util::MbarArriveOnCpAsync(bar);     // fires when this lane's cp.asyncs have landed; bulk bytes are counted by the barrier
            // The announcements of the bulk bytes (expect_tx) and the cp.async-tracked arrivals above reach the barrier
            // through different pipes, in either order (seen in the filter instance: an arrival overtook the expect_tx
            // issued just before it and the phase completed without the tile).  So the phase also waits for one plain
            // arrival that is issued after every producer has issued its announcements: a named barrier among the 256
            // producer threads, then one thread arrives.
            asm volatile("bar.sync 1, %0;" :: "n"(PW * 32) : "memory");
            if (tid == 0) util::MbarArrive(bar);
            i0 = n_i0; i1 = n_i1;
        }
      }
    } else {
        // ================================ consumers ==========================================================
        const int cw = warp - PW;
        unsigned *vmask = ds.d_visited_mask;
        // a consumer reads tile k + 1 into registers (and frees its slot) BEFORE it works on tile k: the slot is back
        // with the producers one tile earlier
        const int qa_i = cw * 64 + lane, qb_i = qa_i + 32;
        int4 qa = make_int4(0, 0, 0, 0), qb = qa;
        auto take = [&](long long k) {
            util::MbarWait(&sm.full[k % N], (unsigned)((k / N) & 1));
            const int4 *t4 = reinterpret_cast<const int4 *>(sm.tile[k % N]);
            qa = t4[qa_i]; qb = t4[qb_i];
            // The slot goes back to the async proxy (its next tenant arrives by TMA): the reads above have to be COMPLETE
            // and ordered before that write.  Issuing them is not enough -- with producers that have nothing else to do
            // (the filter instance) the refill overtook loads still in flight and a few ids per million were read from
            // the next tile instead (found as missing depths at scale 22 / 25).  The fence takes the loaded registers as
            // operands, so it cannot issue before the data is there, and orders the generic reads before the async write.
            asm volatile("fence.proxy.async.shared::cta; // %0 %1 %2 %3 %4 %5 %6 %7"
                         :: "r"(qa.x), "r"(qa.y), "r"(qa.z), "r"(qa.w), "r"(qb.x), "r"(qb.y), "r"(qb.z), "r"(qb.w) : "memory");
            __syncwarp();
            if (lane == 0) util::MbarArrive(&sm.empty[k % N]);          // the ids are in registers: the slot is free
        };
        if (my_tiles > 0) take(0);
        for (long long k = 0; k < my_tiles; ++k) {
            const unsigned long long t = cta + (unsigned long long)k * nctas;
            const ScanT tile_start = t * TILE;
            const int tile_edges = (int)((total - tile_start < (ScanT)TILE) ? (total - tile_start) : (ScanT)TILE);
            if (tid == PW * 32) my_edges += (unsigned long long)tile_edges;
            const VertexId v[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
            if (k + 1 < my_tiles) take(k + 1);
            unsigned valid = 0xffu;
            if (tile_edges < TILE) {                                    // the level's last tile
                valid = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    valid |= (unsigned)(qa_i * 4 + i < tile_edges) << i;
                    valid |= (unsigned)(qb_i * 4 + i < tile_edges) << (4 + i);
                }
            }
            if constexpr (DIRECT) {                                     // holes (-1) a previous operator left in the queue
#pragma unroll
                for (int i = 0; i < 8; ++i) valid &= ~((unsigned)(v[i] < 0) << i);
            }
            // ---- probe: eight words in flight per lane
            unsigned word[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const bool ok = (valid >> i) & 1u, hot = (unsigned)v[i] < hot_bits;
                const unsigned wi = (unsigned)v[i] >> 5;
                word[i] = ProbeVisitedWord(ok && hot, ok && !hot, s_hot_addr + wi * 4u, vmask + wi);
            }
            unsigned live = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) live |= (((word[i] >> (v[i] & 31)) & 1u) ^ 1u) << i;       // ~0u: not valid
            // ---- claim: reductions without a return value, nothing waits for them; HarvestKernel finds the winners
            while (live) {
                const int i = __ffs(live) - 1;
                live &= live - 1;
                const VertexId x = (i & 4) ? ((i & 2) ? ((i & 1) ? v[7] : v[6]) : ((i & 1) ? v[5] : v[4]))
                                           : ((i & 2) ? ((i & 1) ? v[3] : v[2]) : ((i & 1) ? v[1] : v[0]));
                const unsigned bit = 1u << (x & 31), wi = (unsigned)x >> 5;
                if ((unsigned)x < hot_bits) atomicOr(s_hot + wi, bit);  // this SM asks once
                atomicOr(vmask + wi, bit);
            }
        }
    }
    BlockAccumulate<THREADS>(my_edges, sm.red, &slot_out->edges);
}

// The winners of a streamed level: every vertex whose bit is set in the visited mask now and was not in the snapshot
// taken before the level.  Each exactly once and in vertex order.  A warp takes 32 words of the mask at a time (lane <->
// word for the counts, lane <-> vertex of word j for the work), so Functor::ApplyEdge (BFS: the label store), the output
// frontier and the row-offset reads of the out-degree sum are coalesced; one global atomic per 1024 vertices scanned.
template <typename Policy, typename Problem, typename Functor, bool FILTER = false>
__global__ void __launch_bounds__(256)
HarvestKernel(const __grid_constant__ typename Problem::DataSlice ds_in,
              const __grid_constant__ Args<typename Policy::VertexId, typename Policy::SizeT> args,
              const unsigned *before, unsigned long long words)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    constexpr int THREADS = 256;
    typename Problem::DataSlice ds = ds_in;
    if constexpr (HasPartition<typename Problem::DataSlice>::value) ds.part = graph::Partition();
    __shared__ unsigned long long s_red[THREADS / 32 + 1];
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    const int lane = util::LaneId();
    const unsigned below = (1u << lane) - 1;
    const unsigned long long groups = (words + 31) / 32;
    const unsigned long long gw = ((unsigned long long)blockIdx.x * THREADS + threadIdx.x) >> 5;
    const unsigned long long nw = ((unsigned long long)gridDim.x * THREADS) >> 5;
    unsigned long long my_degree = 0;
    for (unsigned long long g = gw; g < groups; g += nw) {
        const unsigned long long w = g * 32 + lane;
        const unsigned bits = (w < words) ? (util::LoadCG(ds.d_visited_mask + w) & ~before[w]) : 0u;
        const unsigned cnt = (unsigned)__popc(bits);
        const unsigned incl = util::WarpInclusiveSum(cnt);
        const unsigned tot = __shfl_sync(util::kFullMask, incl, 31);
        if (tot == 0) continue;                                 // warp-uniform
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&slot_out->out_len, (unsigned long long)tot);
        base = __shfl_sync(util::kFullMask, base, 0);
        for (int j = 0; j < 32; ++j) {
            const unsigned wj = __shfl_sync(util::kFullMask, bits, j);
            const unsigned pj = __shfl_sync(util::kFullMask, incl - cnt, j);
            if (wj == 0) continue;                              // warp-uniform
            if ((wj >> lane) & 1u) {
                const VertexId v = (VertexId)(((g * 32 + j) << 5) + lane);
                if constexpr (FILTER) {
                    // (the streamed filter's queue holds vertices only: CondFilter's "not a hole" test has passed by construction)
                    Functor::ApplyFilter(v, &ds, (typename Problem::Value)args.iteration, (SizeT)0);
                } else {
                    Functor::ApplyEdge((VertexId)args.iteration, v, &ds, (SizeT)0, (SizeT)0);
                }
                if (args.output == OUT_COMPACT) {
                    const unsigned long long at = base + pj + __popc(wj & below);
                    if (at < args.out_cap) args.out_queue[at] = v; else slot_out->overflow = 1;
                }
                if constexpr (Policy::TRACK_OUT_DEGREE)
                    my_degree += (unsigned long long)(args.g.row_offsets[(long long)v + 1] - args.g.row_offsets[v]);
            }
        }
    }
    if constexpr (Policy::TRACK_OUT_DEGREE) BlockAccumulate<THREADS>(my_degree, s_red, &slot_out->out_degree);
}

}  // namespace lb
}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
