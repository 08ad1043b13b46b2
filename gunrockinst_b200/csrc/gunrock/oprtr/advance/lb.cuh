// lb.cuh -- load-balanced ("merge-path") advance for large / skewed frontiers.
//
// Replaces the reference's edge_map_partitioned pipeline
//   GetEdgeCounts -> mgpu::Scan -> cudaMemcpy(total) -> MarkPartitionSizes ->
//   mgpu::SortedSearch -> RelaxPartitionedEdges2 / RelaxLightEdges
// (gunrock/oprtr/advance/kernel.cuh:293-729, oprtr/edge_map_partitioned/
// kernel.cuh:202-559,1172-1463) -- five launches, two library calls and a
// blocking device->host copy per call, 256 CTAs -- by TWO kernels and no host
// round trip:
//
//  1. ScanKernel: degree of every frontier item + single-pass decoupled
//     look-back exclusive scan (own implementation, no CUB / moderngpu), which
//     also records, for every TILE_EDGES-th edge of the expanded frontier, the
//     item that owns it (the merge-path split points; the reference finds them
//     with a separate sorted search).
//  2. ExpandKernel: persistent CTAs (a multiple of the SM count) take
//     equal-sized tiles of the *edge* space.  The items a tile overlaps drop
//     their window index at the tile slot where their list starts and set that
//     slot's bit in a start-flag bitmap (shared memory); thread t handles slots
//     t, t+THREADS, ... so that a warp reads 128 contiguous bytes of a neighbour
//     list, and finds its item with one masked count-leading-zeros on the flag
//     word of its 32 slots (plus a per-word carry every warp derives with one
//     shuffle max-scan) -- no per-edge binary search, 32-bit arithmetic only.
//     Column loads of a whole tile are issued before any of them is consumed,
//     then the optional visited-bitmask probes are issued as a batch, and only
//     survivors reach Functor::CondEdge / ApplyEdge.  Survivors are staged in
//     shared memory by warp-ballot aggregation and flushed to the output
//     frontier with ONE global atomic per ~TILE_EDGES discoveries.
//
// Algorithmic HBM bytes per call: 4 (queue) + 2*sizeof(SizeT) (row offsets) per
// item, 4 per inspected neighbour entry (+1 status byte when PRE_CULL), 4 per
// emitted vertex; scratch traffic (16-24 B per item) is overhead on top.
#pragma once
#include "common.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {
namespace lb {

// ---------------------------------------------------------------------------------
// decoupled look-back tile descriptors: [epoch:14 | status:2 | value:48]
// ---------------------------------------------------------------------------------
constexpr unsigned long long kValueMask  = (1ull << 48) - 1;
constexpr unsigned           kEpochLimit = 1u << 14;
enum { kStatusInvalid = 0, kStatusAggregate = 1, kStatusPrefix = 2 };

__device__ __forceinline__ unsigned long long PackDesc(unsigned epoch, unsigned status, unsigned long long v)
{
    return ((unsigned long long)epoch << 50) | ((unsigned long long)status << 48) | (v & kValueMask);
}
__device__ __forceinline__ unsigned long long LoadDesc(const unsigned long long *p)
{
    return *reinterpret_cast<const volatile unsigned long long *>(p);
}
__device__ __forceinline__ void StoreDesc(unsigned long long *p, unsigned long long v)
{
    *reinterpret_cast<volatile unsigned long long *>(p) = v;
}

// Executed by warp 0 of the CTA that owns `tile`.  Publishes the tile's
// aggregate, walks predecessor descriptors 32 at a time until one carries an
// inclusive prefix, publishes this tile's inclusive prefix and returns its
// exclusive prefix.  Tile numbers are tickets drawn in arrival order, so every
// predecessor tile belongs to a CTA that is already running: the spin cannot
// wait on a CTA that has not been scheduled.
__device__ __forceinline__ unsigned long long LookbackExclusive(
    unsigned long long *desc, unsigned epoch, long long tile, unsigned long long aggregate)
{
    const int lane = util::LaneId();
    if (tile == 0) {
        if (lane == 0) StoreDesc(desc, PackDesc(epoch, kStatusPrefix, aggregate));
        return 0;
    }
    if (lane == 0) StoreDesc(desc + tile, PackDesc(epoch, kStatusAggregate, aggregate));
    unsigned long long exclusive = 0;
    long long look = tile - 1;
    for (;;) {
        long long idx = look - lane;
        unsigned long long d;
        unsigned status;
        do {    // spin until every descriptor of this window has been published this epoch
            d = (idx >= 0) ? LoadDesc(desc + idx) : PackDesc(epoch, kStatusPrefix, 0);
            status = ((unsigned)(d >> 50) == epoch) ? (unsigned)((d >> 48) & 3) : kStatusInvalid;
        } while (__any_sync(util::kFullMask, status == kStatusInvalid));
        unsigned prefix_lanes = __ballot_sync(util::kFullMask, status == kStatusPrefix);
        int first = prefix_lanes ? (__ffs(prefix_lanes) - 1) : 32;
        unsigned long long contrib = (lane <= first) ? (d & kValueMask) : 0ull;
        exclusive += util::WarpReduceSum(contrib);
        if (prefix_lanes) break;
        look -= 32;
    }
    if (lane == 0) StoreDesc(desc + tile, PackDesc(epoch, kStatusPrefix, exclusive + aggregate));
    return exclusive;
}

// ---------------------------------------------------------------------------------
// kernel 1: degrees + exclusive scan + merge-path split points
// ---------------------------------------------------------------------------------
template <int T>
struct ScanStorage {
    unsigned long long warp[T / 32 + 1];
    unsigned long long prefix;
    unsigned           ticket;
};

// Scans tiles of T*R frontier items, handed out by ticket.  `loop` = keep drawing tickets until the
// frontier is exhausted (persistent callers); otherwise exactly one ticket per CTA (the stand-alone
// kernel, whose grid is sized from the host-known length).
template <typename Policy, int T, int R>
__device__ __forceinline__ void ScanTiles(
    const Args<typename Policy::VertexId, typename Policy::SizeT> &args,
    const LbScratchView<typename Policy::SizeT> &scratch,
    ScanStorage<T> &sm, unsigned long long in_len, LevelCounters *slot_out, bool loop)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef unsigned long long        ScanT;
    constexpr int TILE = T * R;
    constexpr unsigned long long TILE_EDGES = Policy::TILE_EDGES;

    do {
        __syncthreads();                    // sm.ticket / sm.prefix of the previous round are consumed
        if (threadIdx.x == 0) sm.ticket = atomicAdd(scratch.ticket, 1u);
        __syncthreads();
        const long long tile = sm.ticket;
        const unsigned long long base = (unsigned long long)tile * TILE;
        if (base >= in_len) {
            // surplus ticket; an empty frontier still has to publish a zero total so that the
            // expand step sees no work
            if (in_len == 0 && tile == 0 && threadIdx.x == 0) { scratch.scanned[0] = 0; slot_out->scan_total = 0; }
            return;
        }

        ScanT deg[R], excl[R];
        SizeT first[R];
        ScanT running = 0;
        // all loads of the tile first (queue, then offsets): the scans below are full of barriers and
        // would otherwise serialise R dependent round trips to memory
        VertexId item[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            unsigned long long i = base + (unsigned long long)r * T + threadIdx.x;
            item[r] = -1;
            if (i < in_len) item[r] = args.in_queue ? util::LoadCG(args.in_queue + i) : (VertexId)args.g.part.Global((long long)i);
        }
        if (InputIsEdges(args.type) && args.in_queue) {         // an edge id stands for the vertex it points to
#pragma unroll
            for (int r = 0; r < R; ++r)
                item[r] = (item[r] >= 0 && (long long)item[r] < args.g.edges) ? args.g.col_indices[item[r]] : (VertexId)-1;
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            deg[r] = 0; first[r] = 0;
            if (item[r] >= 0 && (long long)item[r] < args.g.nodes) {
                long long row = args.g.part.Local(item[r]);
                SizeT a = args.g.row_offsets[row], b = args.g.row_offsets[row + 1];
                first[r] = a; deg[r] = (ScanT)(b - a);
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            ScanT total;
            excl[r] = running + util::BlockExclusiveSum<T>(deg[r], sm.warp, total);
            running += total;
        }
        if (util::WarpId() == 0) {
            ScanT p = LookbackExclusive(scratch.tile_desc, scratch.epoch, tile, running);
            if (util::LaneId() == 0) sm.prefix = p;
        }
        __syncthreads();
        const ScanT prefix = sm.prefix;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            unsigned long long i = base + (unsigned long long)r * T + threadIdx.x;
            if (i < in_len) {
                ScanT b = prefix + excl[r], e = b + deg[r];
                scratch.scanned[i]   = b;
                scratch.row_start[i] = first[r];
                // this item owns every tile boundary that falls inside [b, e)
                for (ScanT k = (b + TILE_EDGES - 1) / TILE_EDGES; k * TILE_EDGES < e && k < scratch.part_cap; ++k)
                    scratch.part_start[k] = (long long)i;
            }
        }
        if (base + TILE >= in_len && threadIdx.x == 0) {
            scratch.scanned[in_len] = prefix + running;
            slot_out->scan_total = prefix + running;
        }
    } while (loop);
}

template <typename Policy>
__global__ void __launch_bounds__(Policy::SCAN_THREADS)
ScanKernel(Args<typename Policy::VertexId, typename Policy::SizeT> args,
           LbScratchView<typename Policy::SizeT> scratch)
{
    __shared__ ScanStorage<Policy::SCAN_THREADS> s_scan;
    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    ScanTiles<Policy, Policy::SCAN_THREADS, Policy::SCAN_ROUNDS>(args, scratch, s_scan, in_len, slot_out, false);
}

// ---------------------------------------------------------------------------------
// kernel 2: tile-wise expansion
// ---------------------------------------------------------------------------------
template <typename Policy, bool NEED_SRC>
struct ExpandStorage {
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    static const int WORDS = Policy::TILE_EDGES / 32;
    int       own[Policy::TILE_EDGES];          // window index of the item whose list starts at this slot
    SizeT     base[Policy::WINDOW];             // CSR position of tile slot y of item j = base[j] + y
    VertexId  src[NEED_SRC ? Policy::WINDOW : 1];
    unsigned  bits[2][WORDS];                   // start-of-list flags, double buffered by tile parity
    StagerStorage<Policy> stager;
    unsigned long long red[Policy::THREADS / 32 + 1];
};

// the vertex behind input item i (an edge id stands for the vertex it points to)
template <typename VertexId, typename SizeT>
__device__ __forceinline__ VertexId SourceOf(const Args<VertexId, SizeT> &args, long long i)
{
    if (!args.in_queue) return (VertexId)args.g.part.Global(i);
    const VertexId x = args.in_queue[i];
    if (InputIsEdges(args.type)) return (x >= 0 && (long long)x < args.g.edges) ? args.g.col_indices[x] : (VertexId)-1;
    return x;
}

// where ExpandTiles finds the scanned frontier: the global scratch ScanKernel filled ...
template <typename SizeT>
struct GlobalScan {
    const LbScratchView<SizeT> &s;
    __device__ __forceinline__ unsigned long long Scanned(long long i) const { return s.scanned[i]; }
    __device__ __forceinline__ SizeT RowStart(long long i) const { return s.row_start[i]; }
    __device__ __forceinline__ long long FirstItem(unsigned long long tile, unsigned long long) const { return s.part_start[tile]; }
    __device__ __forceinline__ unsigned long long TileCap() const { return s.part_cap; }
};
// ... or a CTA-private copy in shared memory (small frontiers: every CTA scans the whole frontier
// itself, which spares the scan pass, its scratch traffic and a grid-wide barrier)
template <typename SizeT>
struct LocalScan {
    const unsigned long long *scanned;      // [in_len + 1]
    const SizeT              *row_start;    // [in_len]
    long long                 in_len;
    __device__ __forceinline__ unsigned long long Scanned(long long i) const { return scanned[i]; }
    __device__ __forceinline__ SizeT RowStart(long long i) const { return row_start[i]; }
    // the item that owns edge tile_start: last i with scanned[i] <= tile_start (it has edges: scanned[i+1] > tile_start)
    __device__ __forceinline__ long long FirstItem(unsigned long long, unsigned long long tile_start) const
    {
        long long lo = 0, hi = in_len + 1;
        while (lo < hi) { const long long mid = (lo + hi) >> 1; if (scanned[mid] <= tile_start) lo = mid + 1; else hi = mid; }
        return lo - 1;
    }
    __device__ __forceinline__ unsigned long long TileCap() const { return ~0ull; }
};

// Expands tiles cta, cta + num_ctas, ... of the scanned frontier.  Callable from any kernel whose
// CTAs have Policy::THREADS threads (the stand-alone ExpandKernel below, the persistent BFS kernel).
template <typename Policy, typename Problem, typename Functor, typename Scan>
__device__ __forceinline__ void ExpandTilesFromImpl(
    const typename Problem::DataSlice &ds,
    const Args<typename Policy::VertexId, typename Policy::SizeT> &args,
    const Scan &scan, const unsigned long long total,
    ExpandStorage<Policy, Problem::MARK_PREDECESSORS> &sm,
    unsigned long long in_len, LevelCounters *slot_out, unsigned cta, unsigned num_ctas)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef unsigned long long        ScanT;
    constexpr int THREADS = Policy::THREADS, EPT = Policy::EDGES_PER_THREAD;
    constexpr int TILE_E = Policy::TILE_EDGES, WINDOW = Policy::WINDOW;
    constexpr int WORDS = TILE_E / 32;
    constexpr bool NEED_SRC = Problem::MARK_PREDECESSORS;
    static_assert(32 % (THREADS / 32) == 0, "warps per CTA must divide 32");

    const int lane = util::LaneId();
    unsigned long long num_tiles = (total + TILE_E - 1) / TILE_E;
    if (num_tiles > scan.TileCap()) {       // a frontier with duplicates expanded past the scratch
        num_tiles = scan.TileCap();
        if (threadIdx.x == 0) slot_out->overflow = 1;
    }
    Stager<Policy> stager(sm.stager, args, slot_out);
    stager.Init();
    for (int x = threadIdx.x; x < 2 * WORDS; x += THREADS) (&sm.bits[0][0])[x] = 0;
    unsigned my_edges = 0;                  // flushed into a 64-bit sum every tile
    unsigned long long my_edges_total = 0;
    int parity = 0;

    for (unsigned long long tile = cta; tile < num_tiles; tile += num_ctas, parity ^= 1) {
        const ScanT tile_start = tile * TILE_E;
        const int tile_edges = (int)((total - tile_start < (ScanT)TILE_E) ? (total - tile_start) : (ScanT)TILE_E);
        const long long i0 = scan.FirstItem(tile, tile_start);
        const long long i1 = (tile + 1 < num_tiles) ? scan.FirstItem(tile + 1, tile_start + TILE_E) : (long long)in_len - 1;
        const int cnt = (int)((i1 - i0 + 1 < (long long)WINDOW + 1) ? (i1 - i0 + 1) : (long long)WINDOW + 1);
        const bool windowed = cnt <= WINDOW;

        __syncthreads();                // previous tile's readers of own/base/src are done; stager count is stable
        if (args.output == OUT_COMPACT) stager.FlushIfNeededSynced(TILE_E);
        unsigned *bits = sm.bits[parity];
        if (windowed) {
            for (int j = threadIdx.x; j < cnt; j += THREADS) {
                const ScanT s = scan.Scanned(i0 + j), e = scan.Scanned(i0 + j + 1);
                const long long rel = (long long)s - (long long)tile_start;     // < 0 only for j == 0
                sm.base[j] = (SizeT)((long long)scan.RowStart(i0 + j) - rel);
                if constexpr (NEED_SRC) sm.src[j] = SourceOf(args, i0 + j);
                const int slot = rel < 0 ? 0 : (int)rel;
                if (e > s && slot < TILE_E) {       // unique: the lists that start inside a tile start at distinct slots
                    sm.own[slot] = j;
                    atomicOr(bits + (slot >> 5), 1u << (slot & 31));
                }
            }
        }
        __syncthreads();
        // item covering the first slot of every 32-slot word: each warp derives all of them with
        // one max-scan (lane L <-> word L of a pass) instead of a CTA-wide step and another barrier
        int carry_reg[(WORDS + 31) / 32];
        if (windowed) {
            int running = 0;
#pragma unroll
            for (int p = 0; p < (WORDS + 31) / 32; ++p) {
                const int wd = p * 32 + lane;
                const unsigned b = (wd < WORDS) ? bits[wd] : 0u;
                int last = b ? sm.own[(wd << 5) + 31 - __clz(b)] : 0;
                last = last > running ? last : running;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int t = __shfl_up_sync(util::kFullMask, last, o);
                    if (lane >= o && t > last) last = t;
                }
                int excl = __shfl_up_sync(util::kFullMask, last, 1);
                if (lane == 0) excl = running;
                carry_reg[p] = excl;
                running = __shfl_sync(util::kFullMask, last, 31);
            }
        }
        // the other flag buffer is free (its readers passed the barrier above): clear it for the next tile
        for (int x = threadIdx.x; x < WORDS; x += THREADS) sm.bits[parity ^ 1][x] = 0;

        SizeT pos[EPT];
        int   item[EPT];
        bool  live[EPT];
#pragma unroll
        for (int r = 0; r < EPT; ++r) {
            const int ly = r * THREADS + threadIdx.x;
            live[r] = ly < tile_edges;
            pos[r] = 0; item[r] = 0;
            if (windowed) {
                const int wd = ly >> 5;                                 // warp-uniform
                const unsigned m = bits[wd] & (0xffffffffu >> (31 - lane));   // starts at or before my slot
                // (wd >> 5 is the same for every warp of the CTA in round r: a compile-time index)
                const int carried = __shfl_sync(util::kFullMask, carry_reg[(r * (THREADS / 32)) >> 5], wd & 31);
                item[r] = m ? sm.own[(wd << 5) + 31 - __clz(m)] : carried;
                pos[r]  = sm.base[item[r]] + (SizeT)ly;
            } else if (live[r]) {                                       // rare: > WINDOW (mostly empty) items in one tile
                const ScanT y = tile_start + ly;
                long long lo = i0, hi = i1 + 1;
                while (lo < hi) { long long mid = (lo + hi) >> 1; if (scan.Scanned(mid) <= y) lo = mid + 1; else hi = mid; }
                item[r] = (int)(lo - 1 - i0);
                pos[r]  = (SizeT)((long long)scan.RowStart(lo - 1) + (long long)(y - scan.Scanned(lo - 1)));
            }
        }
        VertexId nbr[EPT];
#pragma unroll
        for (int r = 0; r < EPT; ++r)
            nbr[r] = live[r] ? util::LoadStream(args.g.col_indices + pos[r]) : (VertexId)-1;
#pragma unroll
        for (int r = 0; r < EPT; ++r) my_edges += live[r] ? 1u : 0u;

        if constexpr (Policy::PRE_CULL) {   // batched visited-bitmask probes (L2 resident)
            unsigned word[EPT];
#pragma unroll
            for (int r = 0; r < EPT; ++r)
                word[r] = live[r] ? util::LoadCA(ds.d_visited_mask + (ds.BitPos(nbr[r]) >> 5)) : 0u;
#pragma unroll
            for (int r = 0; r < EPT; ++r)
                if ((word[r] >> (ds.BitPos(nbr[r]) & 31)) & 1u) live[r] = false;
        }
        // conditions of the thread's entries first, applications afterwards: the conditions' atomics (a
        // visited-bit claim, a distance minimum) are independent and overlap instead of each waiting for
        // the previous entry's answer
        VertexId s_ids[EPT];
        bool     oks[EPT];
#pragma unroll
        for (int r = 0; r < EPT; ++r) {
            oks[r] = false;
            s_ids[r] = (VertexId)args.iteration;
            if (live[r]) {
                if constexpr (NEED_SRC) s_ids[r] = windowed ? sm.src[item[r]] : SourceOf(args, i0 + item[r]);
                oks[r] = Functor::CondEdge(s_ids[r], nbr[r], &ds, pos[r], (SizeT)(i0 + item[r]));
            }
        }
#pragma unroll
        for (int r = 0; r < EPT; ++r) {
            const bool ok = oks[r];
            if (ok) Functor::ApplyEdge(s_ids[r], nbr[r], &ds, pos[r], (SizeT)(i0 + item[r]));
            if (args.output == OUT_COMPACT) {
                stager.Push(ok, OutputIsEdges(args.type) ? (VertexId)pos[r] : nbr[r]);
            } else if (args.output == OUT_TOMBSTONE) {
                const int ly = r * THREADS + threadIdx.x;
                if (ly < tile_edges) {
                    const unsigned long long y = tile_start + ly;
                    if (y < args.out_cap) args.out_queue[y] = ok ? (OutputIsEdges(args.type) ? (VertexId)pos[r] : nbr[r]) : (VertexId)-1;
                    else slot_out->overflow = 1;
                }
            }
        }
        my_edges_total += my_edges; my_edges = 0;
    }
    if (args.output == OUT_COMPACT) stager.Flush();
    else if (args.output == OUT_TOMBSTONE && cta == 0 && threadIdx.x == 0) slot_out->out_len = total;

    // per-CTA statistics: one atomic each
    BlockAccumulate<THREADS>(my_edges_total, sm.red, &slot_out->edges);
    if constexpr (Policy::TRACK_OUT_DEGREE) BlockAccumulate<THREADS>(stager.out_degree, sm.red, &slot_out->out_degree);
}

// does the functor block carry a partition (BC's does not: single device only)?
template <typename T, typename = void> struct HasPartition { static const bool value = false; };
template <typename T> struct HasPartition<T, decltype((void)(((T *)nullptr)->part), void())> { static const bool value = true; };

// Entry: on a single device the functor block is handed over with a compile-time identity partition,
// so the owner hash and the owner-major bit positions drop out of the per-edge code (see pull::Sweep).
template <typename Policy, typename Problem, typename Functor, bool SPECIALISE = true, typename Scan>
__device__ __forceinline__ void ExpandTilesFrom(
    const typename Problem::DataSlice &ds,
    const Args<typename Policy::VertexId, typename Policy::SizeT> &args,
    const Scan &scan, const unsigned long long total,
    ExpandStorage<Policy, Problem::MARK_PREDECESSORS> &sm,
    unsigned long long in_len, LevelCounters *slot_out, unsigned cta, unsigned num_ctas)
{
    if constexpr (SPECIALISE && HasPartition<typename Problem::DataSlice>::value) {
        if (ds.part.log_parts == 0) {
            typename Problem::DataSlice d1 = ds;
            d1.part = graph::Partition();
            ExpandTilesFromImpl<Policy, Problem, Functor>(d1, args, scan, total, sm, in_len, slot_out, cta, num_ctas);
            return;
        }
    }
    ExpandTilesFromImpl<Policy, Problem, Functor>(ds, args, scan, total, sm, in_len, slot_out, cta, num_ctas);
}

// the scanned frontier lives in the global scratch (after ScanKernel / ScanTiles)
template <typename Policy, typename Problem, typename Functor, bool SPECIALISE = true>
__device__ __forceinline__ void ExpandTiles(
    const typename Problem::DataSlice &ds,
    const Args<typename Policy::VertexId, typename Policy::SizeT> &args,
    const LbScratchView<typename Policy::SizeT> &scratch,
    ExpandStorage<Policy, Problem::MARK_PREDECESSORS> &sm,
    unsigned long long in_len, LevelCounters *slot_out, unsigned cta, unsigned num_ctas)
{
    // one load per CTA, not per warp: thousands of warps asking one L2 slice for the same word queue up
    if (threadIdx.x == 0) sm.red[0] = scratch.scanned[in_len];
    __syncthreads();
    const unsigned long long total = sm.red[0];
    GlobalScan<typename Policy::SizeT> scan = {scratch};
    ExpandTilesFrom<Policy, Problem, Functor, SPECIALISE>(ds, args, scan, total, sm, in_len, slot_out, cta, num_ctas);
}

// Small frontiers (in_len <= SMALL_ITEMS): every CTA scans the frontier's degrees itself into shared
// memory and takes its tiles straight away -- one pass, no scratch, no barrier between scan and
// expansion, and a single hub's list is still spread over all CTAs.
template <typename Policy>
struct SmallScanStorage {
    static const int SMALL_ITEMS = Policy::THREADS * 4;
    unsigned long long     scanned[SMALL_ITEMS + 1];
    typename Policy::SizeT row_start[SMALL_ITEMS];
    unsigned long long     warp[Policy::THREADS / 32 + 1];
};

template <typename Policy, typename Problem, typename Functor, bool SPECIALISE = true>
__device__ __forceinline__ void ExpandSmall(
    const typename Problem::DataSlice &ds,
    const Args<typename Policy::VertexId, typename Policy::SizeT> &args,
    SmallScanStorage<Policy> &ss, ExpandStorage<Policy, Problem::MARK_PREDECESSORS> &sm,
    unsigned long long in_len, unsigned long long edges_hint, LevelCounters *slot_out, unsigned cta, unsigned num_ctas)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef unsigned long long        ScanT;
    constexpr int T = Policy::THREADS, R = 4;
    // a CTA beyond the last tile has nothing to do (edges_hint: exact degree sum, or an upper bound)
    if ((unsigned long long)cta * Policy::TILE_EDGES >= edges_hint && cta != 0) return;
    ScanT running = 0;
    VertexId item[R];
    ScanT    deg[R];
    SizeT    first[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {           // all loads first, scans (barriers) afterwards
        const unsigned long long i = (unsigned long long)r * T + threadIdx.x;
        item[r] = -1;
        if (i < in_len) item[r] = args.in_queue ? util::LoadCG(args.in_queue + i) : (VertexId)args.g.part.Global((long long)i);
    }
    if (InputIsEdges(args.type) && args.in_queue) {
#pragma unroll
        for (int r = 0; r < R; ++r)
            item[r] = (item[r] >= 0 && (long long)item[r] < args.g.edges) ? args.g.col_indices[item[r]] : (VertexId)-1;
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        deg[r] = 0; first[r] = 0;
        if (item[r] >= 0 && (long long)item[r] < args.g.nodes) {
            const long long row = args.g.part.Local(item[r]);
            const SizeT a = args.g.row_offsets[row], b = args.g.row_offsets[row + 1];
            first[r] = a; deg[r] = (ScanT)(b - a);
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const unsigned long long i = (unsigned long long)r * T + threadIdx.x;
        if ((unsigned long long)r * T >= in_len) break;         // uniform: no items in this round or later
        ScanT total;
        const ScanT excl = running + util::BlockExclusiveSum<T>(deg[r], ss.warp, total);
        running += total;
        if (i < in_len) { ss.scanned[i] = excl; ss.row_start[i] = first[r]; }
    }
    if (threadIdx.x == 0) ss.scanned[in_len] = running;
    __syncthreads();
    if (cta == 0 && threadIdx.x == 0) slot_out->scan_total = running;
    LocalScan<SizeT> scan = {ss.scanned, ss.row_start, (long long)in_len};
    ExpandTilesFrom<Policy, Problem, Functor, SPECIALISE>(ds, args, scan, running, sm, in_len, slot_out, cta, num_ctas);
}

template <typename Policy, typename Problem, typename Functor>
__global__ void __launch_bounds__(Policy::THREADS, Policy::MIN_CTAS)
ExpandKernel(const __grid_constant__ typename Problem::DataSlice ds,
             const __grid_constant__ Args<typename Policy::VertexId, typename Policy::SizeT> args,
             const __grid_constant__ LbScratchView<typename Policy::SizeT> scratch)
{
    __shared__ ExpandStorage<Policy, Problem::MARK_PREDECESSORS> s_expand;
    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        ClearFutureSlot(args.slots, args.queue_index);
        *scratch.ticket = 0;            // the scan of this call is complete (stream order)
    }
    ExpandTiles<Policy, Problem, Functor>(ds, args, scratch, s_expand, in_len, slot_out, blockIdx.x, gridDim.x);
}

}  // namespace lb
}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
