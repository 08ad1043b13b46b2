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
//     equal-sized tiles of the *edge* space.  A tile stages the scanned degrees
//     of the items it overlaps in shared memory; thread t handles edges
//     t, t+THREADS, ... so that a warp reads 128 contiguous bytes of a neighbour
//     list; each edge finds its item by binary search in shared memory.
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
template <typename Policy>
__global__ void __launch_bounds__(Policy::SCAN_THREADS)
ScanKernel(Args<typename Policy::VertexId, typename Policy::SizeT> args,
           LbScratchView<typename Policy::SizeT> scratch)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef unsigned long long        ScanT;
    constexpr int T = Policy::SCAN_THREADS, R = Policy::SCAN_ROUNDS, TILE = Policy::SCAN_TILE;
    constexpr unsigned long long TILE_EDGES = Policy::TILE_EDGES;

    __shared__ ScanT s_warp[T / 32 + 1];
    __shared__ ScanT s_prefix;
    __shared__ unsigned s_ticket;

    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    if (threadIdx.x == 0) s_ticket = atomicAdd(scratch.ticket, 1u);
    __syncthreads();
    const long long tile = s_ticket;
    const unsigned long long base = (unsigned long long)tile * TILE;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (base >= in_len) {
        // surplus CTA (grid was sized from an upper bound); an empty frontier still
        // has to publish a zero total so that the expand kernel sees no work
        if (in_len == 0 && tile == 0 && threadIdx.x == 0) { scratch.scanned[0] = 0; slot_out->scan_total = 0; }
        return;
    }

    ScanT deg[R], excl[R];
    SizeT first[R];
    ScanT running = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        unsigned long long i = base + (unsigned long long)r * T + threadIdx.x;
        deg[r] = 0; first[r] = 0;
        if (i < in_len) {
            VertexId v = args.in_queue ? util::LoadStream(args.in_queue + i) : (VertexId)args.g.part.Global((long long)i);
            if (v >= 0 && (long long)v < args.g.nodes) {
                long long row = args.g.part.Local(v);
                SizeT a = args.g.row_offsets[row], b = args.g.row_offsets[row + 1];
                first[r] = a; deg[r] = (ScanT)(b - a);
            }
        }
        ScanT total;
        excl[r] = running + util::BlockExclusiveSum<T>(deg[r], s_warp, total);
        running += total;
    }
    if (util::WarpId() == 0) {
        ScanT p = LookbackExclusive(scratch.tile_desc, scratch.epoch, tile, running);
        if (util::LaneId() == 0) s_prefix = p;
    }
    __syncthreads();
    const ScanT prefix = s_prefix;
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
}

// ---------------------------------------------------------------------------------
// kernel 2: tile-wise expansion
// ---------------------------------------------------------------------------------
template <typename Policy, typename Problem, typename Functor>
__global__ void __launch_bounds__(Policy::THREADS, Policy::MIN_CTAS)
ExpandKernel(const __grid_constant__ typename Problem::DataSlice ds,
             const __grid_constant__ Args<typename Policy::VertexId, typename Policy::SizeT> args,
             const __grid_constant__ LbScratchView<typename Policy::SizeT> scratch)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    typedef unsigned long long        ScanT;
    constexpr int THREADS = Policy::THREADS, EPT = Policy::EDGES_PER_THREAD;
    constexpr int TILE_E = Policy::TILE_EDGES, WINDOW = Policy::WINDOW, STAGE_CAP = Policy::STAGE_CAP;
    constexpr bool NEED_SRC = Problem::MARK_PREDECESSORS;

    __shared__ int        s_rel[WINDOW];        // scanned degree of item j minus tile start
    __shared__ long long  s_base[WINDOW];       // CSR position of edge y of item j = s_base[j] + y
    __shared__ VertexId   s_src[NEED_SRC ? WINDOW : 1];
    __shared__ StagerStorage<Policy> s_stager;
    __shared__ unsigned long long s_red[THREADS / 32 + 1];

    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        ClearFutureSlot(args.slots, args.queue_index);
        *scratch.ticket = 0;            // the scan of this call is complete (stream order)
    }

    const ScanT total = scratch.scanned[in_len];
    unsigned long long num_tiles = (total + TILE_E - 1) / TILE_E;
    if (num_tiles > scratch.part_cap) {     // a frontier with duplicates expanded past the scratch
        num_tiles = scratch.part_cap;
        if (threadIdx.x == 0) slot_out->overflow = 1;
    }
    Stager<Policy> stager(s_stager, args, slot_out);
    stager.Init();
    unsigned long long my_edges = 0;

    for (unsigned long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const ScanT tile_start = tile * TILE_E;
        const int tile_edges = (int)((total - tile_start < (ScanT)TILE_E) ? (total - tile_start) : (ScanT)TILE_E);
        const long long i0 = scratch.part_start[tile];
        const long long i1 = (tile + 1 < num_tiles) ? scratch.part_start[tile + 1] : (long long)in_len - 1;
        const long long cnt = i1 - i0 + 1;
        const bool windowed = cnt <= WINDOW;

        __syncthreads();                // previous tile's readers of s_rel/s_base are done
        if (windowed) {
            for (int j = threadIdx.x; j < (int)cnt; j += THREADS) {
                ScanT s = scratch.scanned[i0 + j];
                long long rel = (long long)s - (long long)tile_start;
                s_rel[j]  = rel < -1 ? -1 : (int)rel;
                s_base[j] = (long long)scratch.row_start[i0 + j] - (long long)s;
                if constexpr (NEED_SRC) s_src[j] = args.in_queue ? args.in_queue[i0 + j] : (VertexId)args.g.part.Global(i0 + j);
            }
        }
        __syncthreads();

        long long pos[EPT];             // CSR entry index, -1 = no edge for this slot
        int       item[EPT];
#pragma unroll
        for (int r = 0; r < EPT; ++r) {
            const int ly = r * THREADS + threadIdx.x;
            pos[r] = -1; item[r] = 0;
            if (ly < tile_edges) {
                if (windowed) {
                    int lo = 0, hi = (int)cnt;          // first j with s_rel[j] > ly
                    while (lo < hi) { int mid = (lo + hi) >> 1; if (s_rel[mid] <= ly) lo = mid + 1; else hi = mid; }
                    item[r] = lo - 1;
                    pos[r]  = s_base[lo - 1] + (long long)(tile_start + ly);
                } else {                                 // rare: > WINDOW (mostly empty) items in one tile
                    const ScanT y = tile_start + ly;
                    long long lo = i0, hi = i1 + 1;
                    while (lo < hi) { long long mid = (lo + hi) >> 1; if (scratch.scanned[mid] <= y) lo = mid + 1; else hi = mid; }
                    item[r] = (int)(lo - 1 - i0);
                    pos[r]  = (long long)scratch.row_start[lo - 1] + (long long)(y - scratch.scanned[lo - 1]);
                }
            }
        }
        VertexId nbr[EPT];
#pragma unroll
        for (int r = 0; r < EPT; ++r)
            nbr[r] = (pos[r] >= 0) ? util::LoadStream(args.g.col_indices + pos[r]) : (VertexId)-1;

        bool live[EPT];
#pragma unroll
        for (int r = 0; r < EPT; ++r) {
            live[r] = pos[r] >= 0;
            if (live[r]) ++my_edges;
        }
        if constexpr (Policy::PRE_CULL) {   // batched visited-bitmask probes (L2 resident)
            unsigned word[EPT];
#pragma unroll
            for (int r = 0; r < EPT; ++r)
                word[r] = live[r] ? util::LoadCA(ds.d_visited_mask + (ds.BitPos(nbr[r]) >> 5)) : 0u;
#pragma unroll
            for (int r = 0; r < EPT; ++r)
                if (live[r] && ((word[r] >> (ds.BitPos(nbr[r]) & 31)) & 1u)) live[r] = false;
        }
#pragma unroll
        for (int r = 0; r < EPT; ++r) {
            bool ok = false;
            if (live[r]) {
                VertexId s_id;
                if constexpr (NEED_SRC) s_id = windowed ? s_src[item[r]]
                                      : (args.in_queue ? args.in_queue[i0 + item[r]] : (VertexId)args.g.part.Global(i0 + item[r]));
                else                    s_id = (VertexId)args.iteration;
                ok = VisitEdge<Policy, Problem, Functor>(&ds, s_id, nbr[r], (SizeT)pos[r], (SizeT)(i0 + item[r]));
            }
            if (args.output == OUT_COMPACT) {
                stager.Push(ok, nbr[r]);
            } else if (args.output == OUT_TOMBSTONE) {
                const int ly = r * THREADS + threadIdx.x;
                if (ly < tile_edges) {
                    const unsigned long long y = tile_start + ly;
                    if (y < args.out_cap) args.out_queue[y] = ok ? nbr[r] : (VertexId)-1;
                    else slot_out->overflow = 1;
                }
            }
        }
        if (args.output == OUT_COMPACT) stager.FlushIfNeeded(TILE_E);
    }
    if (args.output == OUT_COMPACT) stager.Flush();
    else if (args.output == OUT_TOMBSTONE && blockIdx.x == 0 && threadIdx.x == 0) slot_out->out_len = total;

    // per-CTA statistics: one atomic each
    BlockAccumulate<THREADS>(my_edges, s_red, &slot_out->edges);
    if constexpr (Policy::TRACK_OUT_DEGREE) BlockAccumulate<THREADS>(stager.out_degree, s_red, &slot_out->out_degree);
}

}  // namespace lb
}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
