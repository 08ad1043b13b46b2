// pull.cuh -- pull-direction ("bottom-up", backward) advance for direction-optimising BFS.
//
// Replaces the reference's edge_map_backward / edge_map_partitioned_backward
// (gunrock/oprtr/edge_map_backward/cta.cuh:220-341,652-773;
// oprtr/edge_map_partitioned_backward/kernel.cuh:377-526), which keep the
// frontier as `bool` BYTES (dobfs_problem.cuh:63-64), materialise an
// unvisited-vertex queue from an iota array with two extra filter passes
// (dobfs_enactor.cuh:422-462) and have no cross-thread early exit.
//
// Here the state is three 1-bit-per-vertex maps (n/8 bytes each; 4 MiB at scale
// 25, resident in the 126 MB L2):
//     visited        read+written, only the words of the rows this device owns
//     frontier_in    read-only (probed through the non-coherent path)
//     frontier_out   written whole words, no atomics, no prior memset
// and the kernel SWEEPS the owned rows in chunks of CHUNK_WORDS bitmap words
// (256 vertices) per warp:
//   1. scan    -- the warp reads the chunk's visited / candidate words and packs the
//                 positions of the vertices still to do into a shared-memory list
//                 (byte offsets), so that the walk below runs with dense lanes
//                 whatever fraction of the chunk is left (late pull levels leave < 1 %);
//                 a chunk with nothing to do costs two coalesced loads per word;
//   1'. dense  -- a chunk with at least kDenseMin vertices to do (the first sweeps of a search) skips the
//                 list: lane l takes bit l of each word, so the heads, offsets and label stores of a word are
//                 coalesced accesses of 32 consecutive rows, a word's discoveries are one ballot, and only
//                 the vertices whose first entry was no parent are listed for the walk (DenseChunk);
//   2. probe   -- phase-wise over the WHOLE list (up to 8 vertices per lane): first all
//                 row-offset pairs, then all first-neighbour ids, then all frontier-bit
//                 probes, so the dependent chain offsets -> neighbour -> bit is three
//                 loads deep for the chunk, not per vertex, with up to 8 independent
//                 loads in flight per lane in every phase (an R-MAT list is sorted, hubs
//                 first, so ~90 % of the vertices find their parent at their first entry);
//   3. walk    -- the vertices that did not are packed into a retry list and continue
//                 lane-serially with early exit up to SERIAL_LIMIT entries; what is left
//                 of long lists is finished by the whole warp with 128-byte coalesced
//                 reads and a ballot early exit, so a straggler list cannot serialise a warp;
//   4. publish -- found vertices are OR-ed into a per-warp word buffer in shared memory
//                 and leave as whole-word stores (frontier_out, visited), no global atomics.
//
// Optional `head` array (first entry of every row, 4 bytes per vertex, built once per graph):
// a sorted R-MAT list starts with its lowest-numbered, i.e. highest-degree, neighbour, so nine
// vertices out of ten find their parent there; reading that entry from a dense array instead
// of col_indices[row_offsets[v]] replaces one 128-byte line (and one 32-byte DRAM sector) per
// vertex by coalesced reads of consecutive rows.
//
// Algorithmic HBM bytes per call: n/8 (visited) [+ n/8 candidates] + n/8
// (frontier_out) for the sweep, 2*sizeof(SizeT) per unvisited candidate, 4 + 1
// per inspected neighbour entry (column id + frontier bit probe), 4 per found
// vertex (label).
#pragma once
#ifndef GRB_PULL_OVERLAP
#define GRB_PULL_OVERLAP 0
#endif
#include "common.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {
namespace pull {

template <typename VertexId, typename SizeT>
struct PullArgs {
    graph::CsrView<VertexId, SizeT> g;      // INVERSE graph (CSC); == CSR when symmetric
    const unsigned *frontier_in;            // global bitmap, owner-major bit positions
    unsigned       *frontier_out;           // global bitmap; this device writes its own words
    unsigned       *visited;                // global bitmap; this device updates its own words
    const unsigned *candidates;             // optional, local words: 1 = vertex has in-edges
    const VertexId *head;                   // optional, per local row: first entry of its list (-1: empty)
    unsigned long long *tickets;            // optional: kTicketCounters zeroed counters, kTicketStride words apart
                                            // (dynamic chunk hand-out; NULL: static stride)
    const SizeT    *fwd_row_offsets;        // forward (out-edge) offsets for the degree sum; NULL: use g
    VertexId       *out_queue;              // optional: also enqueue found vertices
    unsigned long long out_cap;
    // partitioned runs with peer-memory exchange: the same frontier map on the other devices
    // (NVLink mappings).  Every word this device publishes is also stored there, so the map is
    // complete everywhere when the level ends and no all-gather pass follows the sweep.
    unsigned       *peer_out[7] = {};
    int             peers = 0;
    LevelCounters  *slots;
    long long       queue_index;
    long long       iteration;
};


constexpr int kChunkWords = 8;                  // bitmap words a warp sweeps at a time
constexpr int kChunkVerts = kChunkWords * 32;   // ... = 256 vertices, so list entries fit a byte
constexpr int kChunkPerLane = kChunkVerts / 32;
#ifndef GRB_PULL_DENSE_MIN
#define GRB_PULL_DENSE_MIN 48
#endif
#ifndef GRB_PULL_WALK_BATCH
#define GRB_PULL_WALK_BATCH 4
#endif
#ifndef GRB_PULL_DENSE_GROUP
#define GRB_PULL_DENSE_GROUP 4
#endif
constexpr int kDenseMin = GRB_PULL_DENSE_MIN;   // a chunk with at least this many vertices to do takes the dense path (> 256: never)
constexpr int kTicketCounters = 16;             // chunk tickets are drawn from 16 counters on 16 different L2 lines:
constexpr int kTicketStride   = 16;             // one counter would serialise ~16K atomics per sweep at scale 22

template <typename Policy>
struct SweepStorage {
    typedef typename Policy::SizeT SizeT;
    SizeT              begin[Policy::WARPS][kChunkVerts];   // per listed vertex: first entry of its list ...
    unsigned           deg[Policy::WARPS][kChunkVerts];     // ... and its length (clamped to 2^31 - 1)
    unsigned char      cand[Policy::WARPS][kChunkVerts];    // per warp: positions of the chunk's vertices still to do
    unsigned char      retry[Policy::WARPS][kChunkVerts];   // list slots whose first entry was not in the frontier
    unsigned           found[Policy::WARPS][kChunkWords];   // per warp: discoveries of the chunk, one bit each
    unsigned long long red[Policy::WARPS + 1];
};

// what a dense chunk adds to the warp's running sums (see SweepImpl)
struct DenseResult {
    unsigned n_retry, cands, edges;
    unsigned long long degree;
};

// The dense-chunk path of the sweep (steps 1' + 2' in SweepImpl): lane l takes bit l of each of the chunk's words.
// (Measured as a __noinline__ function, to keep its registers away from the list path: 45 % slower, scale 22 and 25.)
template <typename Policy, typename Problem, typename Functor, bool SINGLE>
__device__ __forceinline__ DenseResult DenseChunk(
    const typename Problem::DataSlice &ds,
    const PullArgs<typename Policy::VertexId, typename Policy::SizeT> &args,
    LevelCounters *slot_out, long long w0, unsigned todo,
    unsigned char *s_cand, unsigned char *s_retry, typename Policy::SizeT *s_begin, unsigned *s_deg, unsigned *s_found)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    constexpr bool NEED_SRC = Problem::MARK_PREDECESSORS;
    const int lane = util::LaneId();
    DenseResult r;
    r.n_retry = 0; r.cands = 0; r.edges = 0; r.degree = 0;
    {
            constexpr int G = GRB_PULL_DENSE_GROUP;                                // words per group: G independent loads per phase
#pragma unroll
            for (int g = 0; g < kChunkWords; g += G) {
                VertexId h[G];
                SizeT    rb[G], re[G];
                unsigned pr[G];
#pragma unroll
                for (int j = 0; j < G; ++j) {
                    const unsigned tw = __shfl_sync(util::kFullMask, todo, (g + j) << 2);
                    const long long row = ((w0 + g + j) << 5) + lane;
                    h[j] = -1; rb[j] = 0; re[j] = 0;
                    if ((tw >> lane) & 1u) {
                        h[j]  = util::LoadStream(args.head + row);
                        rb[j] = args.g.row_offsets[row];
                        re[j] = args.g.row_offsets[row + 1];
                        ++r.cands;
                    }
                }
#pragma unroll
                for (int j = 0; j < G; ++j)
                    pr[j] = h[j] >= 0 ? util::LoadCA(args.frontier_in + (ds.BitPos(h[j]) >> 5)) : 0u;
#pragma unroll
                for (int j = 0; j < G; ++j) {
                    const bool hit = h[j] >= 0 && ((pr[j] >> (ds.BitPos(h[j]) & 31)) & 1u);
                    const unsigned long long d = (unsigned long long)(re[j] - rb[j]);
                    bool ok = false;
                    if (h[j] >= 0) ++r.edges;
                    if (hit) {
                        const long long row = ((w0 + g + j) << 5) + lane;
                        const VertexId child = SINGLE ? (VertexId)row : (VertexId)args.g.part.Global(row);
                        const VertexId s_id = NEED_SRC ? h[j] : (VertexId)args.iteration;
                        ok = VisitEdge<Policy, Problem, Functor>(&ds, s_id, child, rb[j], (SizeT)row);
                        if (ok) {
                            if (args.fwd_row_offsets)
                                r.degree += (unsigned long long)(args.fwd_row_offsets[row + 1] - args.fwd_row_offsets[row]);
                            else
                                r.degree += d;
                            if (args.out_queue) {
                                const unsigned long long at = atomicAdd(&slot_out->out_len, 1ull);
                                if (at < args.out_cap) args.out_queue[at] = child;
                                else slot_out->overflow = 1;
                            }
                        }
                    }
                    const unsigned fm = __ballot_sync(util::kFullMask, ok);
                    if (lane == 0) s_found[g + j] = fm;
                    const bool again = !hit && d > 1;
                    const unsigned am = __ballot_sync(util::kFullMask, again);
                    if (again) {
                        const unsigned p = r.n_retry + __popc(am & ((1u << lane) - 1));
                        s_cand[p]  = (unsigned char)((g + j) * 32 + lane);
                        s_begin[p] = rb[j];
                        s_deg[p]   = d > 0x7fffffffull ? 0x7fffffffu : (unsigned)d;
                        s_retry[p] = (unsigned char)p;
                    }
                    r.n_retry += (unsigned)__popc(am);
                }
            }
    }
    return r;
}

// The sweep itself, callable from any kernel whose CTAs have Policy::THREADS threads:
// warp `first_chunk` of `chunk_stride` warps takes chunks first_chunk, first_chunk + chunk_stride, ...
// Accumulates edges / out_degree / aux0 (= vertices examined) / out_len into `slot_out`.
template <typename Policy, typename Problem, typename Functor, bool SINGLE>
__device__ __forceinline__ void SweepImpl(
    const typename Problem::DataSlice &ds,
    const PullArgs<typename Policy::VertexId, typename Policy::SizeT> &args,
    SweepStorage<Policy> &sm, LevelCounters *slot_out,
    long long first_chunk, long long chunk_stride)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    constexpr int THREADS = Policy::THREADS;
    constexpr int SERIAL_LIMIT = 8, U = kChunkPerLane;
    static_assert(sizeof(VertexId) == 4, "the sweep fetches neighbour ids as aligned groups of four 32-bit words");
    constexpr bool NEED_SRC = Problem::MARK_PREDECESSORS;

    const int lane = util::LaneId();
    const long long local_nodes = args.g.local_nodes;
    const long long num_words  = (local_nodes + 31) >> 5;
    const long long num_chunks = (num_words + kChunkWords - 1) / kChunkWords;
    // first word of this device's segment inside the global bitmaps
    const long long seg = SINGLE ? 0ll : ((long long)args.g.part.rank * local_nodes) >> 5;
    unsigned char *s_cand  = sm.cand[util::WarpId()];
    unsigned char *s_retry = sm.retry[util::WarpId()];
    SizeT         *s_begin = sm.begin[util::WarpId()];
    unsigned      *s_deg   = sm.deg[util::WarpId()];
    unsigned      *s_found = sm.found[util::WarpId()];
    const int my_word = lane >> 2, my_shift = (lane & 3) << 3;     // lane -> byte (lane & 3) of word (lane >> 2)

    unsigned long long my_edges = 0, my_found = 0, my_degree = 0, my_cands = 0;

    // functor + publication of one found vertex (list slot idx of the current chunk)
    auto publish = [&](long long w0, unsigned idx, VertexId parent, long long epos) {
        const unsigned off = s_cand[idx];
        const long long row = (w0 << 5) + off;
        const VertexId child = SINGLE ? (VertexId)row : (VertexId)args.g.part.Global(row);
        const VertexId s_id = NEED_SRC ? parent : (VertexId)args.iteration;
        if (!VisitEdge<Policy, Problem, Functor>(&ds, s_id, child, (SizeT)epos, (SizeT)row)) return;
        atomicOr(s_found + (off >> 5), 1u << (off & 31));
        if (args.fwd_row_offsets)
            my_degree += (unsigned long long)(args.fwd_row_offsets[row + 1] - args.fwd_row_offsets[row]);
        else
            my_degree += s_deg[idx];
        if (args.out_queue) {
            const unsigned long long at = atomicAdd(&slot_out->out_len, 1ull);
            if (at < args.out_cap) args.out_queue[at] = child;
            else slot_out->overflow = 1;
        }
    };

    // chunks: the first one by position, the following ones by ticket (chunks differ a lot in cost and
    // 16384 chunks over 4736 resident warps is 3 or 4 each: a static stride left 15 % of the warp time
    // waiting at the final barrier); the ticket for the next chunk is drawn before the current one is
    // processed, so its round trip is hidden
    long long c = first_chunk;
    const int my_counter = (int)(first_chunk & (kTicketCounters - 1));
    unsigned long long *counter = args.tickets ? args.tickets + my_counter * kTicketStride : nullptr;
    while (c < num_chunks) {
        unsigned long long ticket = 0;
        if (lane == 0 && counter) ticket = atomicAdd(counter, 1ull);
        // ---- 1. scan ---------------------------------------------------------------------------
        const long long w0 = c * kChunkWords, w = w0 + my_word;
        unsigned vis = 0xffffffffu, todo = 0;
        if (w < num_words) {
            vis  = util::LoadCG(args.visited + seg + w);
            todo = ~vis;
            if (args.candidates) todo &= util::LoadStream(args.candidates + w);
            if (w == num_words - 1 && (local_nodes & 31)) todo &= (1u << (local_nodes & 31)) - 1;
        }
        const unsigned mine = (todo >> my_shift) & 0xffu;
        const unsigned cnt  = (unsigned)__popc(mine);
        const unsigned incl = util::WarpInclusiveSum(cnt);
        const unsigned total = __shfl_sync(util::kFullMask, incl, 31);
        if (total == 0) {                                       // uniform across the warp
            if ((lane & 3) == 0 && w < num_words) {
                args.frontier_out[seg + w] = 0;
                for (int p = 0; p < args.peers; ++p) args.peer_out[p][seg + w] = 0;
            }
            c = counter ? chunk_stride + (long long)__shfl_sync(util::kFullMask, ticket, 0) * kTicketCounters + my_counter
                        : c + chunk_stride;
            continue;
        }
        unsigned n_retry = 0;
        if (kDenseMin <= kChunkVerts && args.head != nullptr && total >= (unsigned)kDenseMin) {
            // ---- 1' + 2'. dense chunk: lane l takes bit l of each word, no candidate list --------------
            // (early pull levels: most of the chunk is still to do, so packing positions into a list
            //  buys nothing and its bookkeeping -- ~11 warp instructions per candidate -- was what bound
            //  the sweep; here heads, offsets and label stores of a word are coalesced accesses of 32
            //  consecutive rows, the discoveries of a word are one ballot, and only the misses are listed)
            const DenseResult r = DenseChunk<Policy, Problem, Functor, SINGLE>(ds, args, slot_out, w0, todo, s_cand, s_retry,
                                                                                   s_begin, s_deg, s_found);
            n_retry = r.n_retry; my_cands += r.cands; my_edges += r.edges; my_degree += r.degree;
        } else {
        {
            unsigned at = incl - cnt, m = mine;
            const unsigned pos0 = (unsigned)(my_word * 32 + my_shift);
            while (m) { const int b = __ffs(m) - 1; m &= m - 1; s_cand[at++] = (unsigned char)(pos0 + b); }
            if (lane < kChunkWords) s_found[lane] = 0;
        }
        __syncwarp();
        my_cands += cnt;

        // ---- 2. probe, phase by phase over the whole list -------------------------------------------
        // (u * 32 < total is warp-uniform: unused phases of a short list are skipped)
#if GRB_PULL_OVERLAP
        // the dense head array (first entry of every row, -1: empty row) is indexed by the row like the
        // offsets, so both gathers are issued together and the frontier probes below wait for the
        // head only: the chain is head -> bit, with the offsets (needed for the degree sum and the
        // retry walk) arriving in its shadow
        VertexId n0[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned idx = u * 32 + lane;
            n0[u] = -1;
            if (args.head && u * 32 < total && idx < total)
                n0[u] = util::LoadStream(args.head + ((w0 << 5) + s_cand[idx]));
        }
        SizeT rb[U], re[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned idx = u * 32 + lane;
            rb[u] = 0; re[u] = 0;
            if (u * 32 < total && idx < total) {
                const long long row = (w0 << 5) + s_cand[idx];
                rb[u] = args.g.row_offsets[row]; re[u] = args.g.row_offsets[row + 1];
            }
        }
        if (!args.head) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const unsigned idx = u * 32 + lane;
                if (u * 32 < total && idx < total && re[u] > rb[u])
                    n0[u] = util::LoadStream(args.g.col_indices + rb[u]);
            }
        }
        unsigned probe[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            probe[u] = 0;
            if (u * 32 < total && n0[u] >= 0) probe[u] = util::LoadCA(args.frontier_in + (ds.BitPos(n0[u]) >> 5));
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned idx = u * 32 + lane;
            if (u * 32 < total && idx < total) {
                s_begin[idx] = rb[u];
                const unsigned long long d = (unsigned long long)(re[u] - rb[u]);
                s_deg[idx] = d > 0x7fffffffull ? 0x7fffffffu : (unsigned)d;
            }
        }
        unsigned hits = 0;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (u * 32 < total && n0[u] >= 0) {
                ++my_edges;
                if ((probe[u] >> (ds.BitPos(n0[u]) & 31)) & 1u) hits |= 1u << u;
            }
        }
#else
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned idx = u * 32 + lane;
            if (u * 32 < total && idx < total) {
                const long long row = (w0 << 5) + s_cand[idx];
                const SizeT b = args.g.row_offsets[row], e = args.g.row_offsets[row + 1];
                s_begin[idx] = b;
                const unsigned long long d = (unsigned long long)(e - b);
                s_deg[idx] = d > 0x7fffffffull ? 0x7fffffffu : (unsigned)d;
            }
        }
        VertexId n0[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const unsigned idx = u * 32 + lane;
            n0[u] = -1;
            if (u * 32 < total && idx < total && s_deg[idx] != 0) {
                // the dense head array turns this gather (one 128-byte line per vertex) into a
                // coalesced read of consecutive rows
                if (args.head) n0[u] = util::LoadStream(args.head + ((w0 << 5) + s_cand[idx]));
                else           n0[u] = util::LoadStream(args.g.col_indices + s_begin[idx]);
            }
        }
        unsigned hits = 0;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (u * 32 < total && n0[u] >= 0) {
                ++my_edges;
                if (util::BitTest(args.frontier_in, ds.BitPos(n0[u]))) hits |= 1u << u;
            }
        }
#endif
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (u * 32 < total) {                               // uniform
                const unsigned idx = u * 32 + lane;
                const bool hit = (hits >> u) & 1u;
                if (hit) publish(w0, idx, n0[u], (long long)s_begin[idx]);
                const bool again = !hit && idx < total && s_deg[idx] > 1;
                const unsigned am = __ballot_sync(util::kFullMask, again);
                if (again) s_retry[n_retry + __popc(am & ((1u << lane) - 1))] = (unsigned char)idx;
                n_retry += (unsigned)__popc(am);
            }
        }
        }   // list path
        __syncwarp();
        // ---- 3. walk what is left ------------------------------------------------------------------
        for (unsigned k0 = 0; k0 < n_retry; k0 += 32) {
            const bool act = k0 + lane < n_retry;
            const unsigned idx = act ? (unsigned)s_retry[k0 + lane] : 0u;
            long long begin = 0, end = 0, k = 0, epos = 0;
            bool found = false;
            VertexId parent = -1;
            if (act) {
                begin = (long long)s_begin[idx];
                end = begin + (long long)s_deg[idx];
                if (s_deg[idx] == 0x7fffffffu) {                // clamped: re-read the true end
                    const long long row = (w0 << 5) + s_cand[idx];
                    end = (long long)args.g.row_offsets[row + 1];
                }
                k = begin + 1;
                const long long lim = (end - begin > SERIAL_LIMIT) ? begin + SERIAL_LIMIT : end;
#if GRB_PULL_WALK_BATCH > 1
                // WB entries and their frontier words are requested together: the chain of a list shrinks from
                // two dependent loads per entry to two per batch (a sweep with 3-4 chunks per warp is bound by the
                // latency of its slowest lane, not by throughput); the entries counted are those an early-exit
                // walk looks at, as before.  Measured with the dense-chunk path (profiles/r02_pull_dense_ab.txt):
                // WB = 4 takes the first sweep of a scale-22 search from 48 to 41 us, WB = 8 spills and loses it
                // again; without the dense path (every vertex walks through the list) batches were no gain.
                constexpr int WB = GRB_PULL_WALK_BATCH;
                while (k < lim) {
                    VertexId x[WB];
                    unsigned pw[WB];
#pragma unroll
                    for (int b = 0; b < WB; ++b) x[b] = (k + b < lim) ? util::LoadStream(args.g.col_indices + k + b) : (VertexId)-1;
#pragma unroll
                    for (int b = 0; b < WB; ++b) pw[b] = x[b] >= 0 ? util::LoadCA(args.frontier_in + (ds.BitPos(x[b]) >> 5)) : 0u;
                    int first = -1;
#pragma unroll
                    for (int b = WB - 1; b >= 0; --b)
                        if (x[b] >= 0 && ((pw[b] >> (ds.BitPos(x[b]) & 31)) & 1u)) { first = b; parent = x[b]; }
                    if (first >= 0) { found = true; epos = k + first; k = epos + 1; break; }
                    k = (k + WB < lim) ? k + WB : lim;
                }
#else
                while (k < lim) {
                    const VertexId x = util::LoadStream(args.g.col_indices + k);
                    if (util::BitTest(args.frontier_in, ds.BitPos(x))) { found = true; parent = x; epos = k; ++k; break; }
                    ++k;
                }
#endif
                my_edges += (unsigned long long)(k - begin - 1);    // entries an early-exit walk looks at
            }
            unsigned rest = __ballot_sync(util::kFullMask, act && !found && k < end);
            while (rest) {
                const int leader = __ffs(rest) - 1;
                rest &= rest - 1;
                const long long wb = __shfl_sync(util::kFullMask, k, leader);
                const long long we = __shfl_sync(util::kFullMask, end, leader);
                for (long long p = wb; p < we; p += 32) {
                    const long long at = p + lane;
                    VertexId x = -1;
                    bool hit = false;
                    if (at < we) {
                        x = util::LoadStream(args.g.col_indices + at);
                        hit = util::BitTest(args.frontier_in, ds.BitPos(x));
                        ++my_edges;
                    }
                    const unsigned hm = __ballot_sync(util::kFullMask, hit);
                    if (hm) {
                        const int first = __ffs(hm) - 1;
                        const VertexId px = __shfl_sync(util::kFullMask, x, first);
                        if (lane == leader) { found = true; parent = px; epos = p + first; }
                        break;
                    }
                }
            }
            if (found) publish(w0, idx, parent, epos);
        }
        __syncwarp();
        // ---- 4. whole-word publication: lane l < CHUNK_WORDS owns word w0 + l (its visited word sits in lane 4l)
        const unsigned vis_l = __shfl_sync(util::kFullMask, vis, (lane & (kChunkWords - 1)) << 2);
        if (lane < kChunkWords && w0 + lane < num_words) {
            const unsigned fm = s_found[lane];
            args.frontier_out[seg + w0 + lane] = fm;
            for (int p = 0; p < args.peers; ++p) args.peer_out[p][seg + w0 + lane] = fm;
            if (fm) args.visited[seg + w0 + lane] = vis_l | fm;
            my_found += (unsigned)__popc(fm);
        }
        __syncwarp();                                           // the lists are rewritten by the next chunk
        c = counter ? chunk_stride + (long long)__shfl_sync(util::kFullMask, ticket, 0) * kTicketCounters + my_counter
                    : c + chunk_stride;
    }
    BlockAccumulate<THREADS>(my_edges,  sm.red, &slot_out->edges);
    BlockAccumulate<THREADS>(my_degree, sm.red, &slot_out->out_degree);
    BlockAccumulate<THREADS>(my_cands,  sm.red, &slot_out->aux0);
    // found vertices are counted in out_len unless the queue path already did
    if (!args.out_queue) BlockAccumulate<THREADS>(my_found, sm.red, &slot_out->out_len);
}

// Entry: on a single device the partition arithmetic (owner hash, owner-major bit positions, global
// <-> local ids) is the identity.  Handing the implementation copies whose partition is a
// compile-time constant lets the compiler drop it from every per-vertex and per-edge path (it was
// ~10 % of the sweep's instructions); the partitioned run takes the general instance.
template <typename Policy, typename Problem, typename Functor, bool SPECIALISE = true>
__device__ __forceinline__ void Sweep(
    const typename Problem::DataSlice &ds,
    const PullArgs<typename Policy::VertexId, typename Policy::SizeT> &args,
    SweepStorage<Policy> &sm, LevelCounters *slot_out,
    long long first_chunk, long long chunk_stride)
{
    if (SPECIALISE && args.g.part.log_parts == 0) {
        typename Problem::DataSlice d1 = ds;       // (only the small functor block is copied: a copy
        d1.part = graph::Partition();               //  of `args` ends up in local memory)
        SweepImpl<Policy, Problem, Functor, true>(d1, args, sm, slot_out, first_chunk, chunk_stride);
    } else {
        SweepImpl<Policy, Problem, Functor, false>(ds, args, sm, slot_out, first_chunk, chunk_stride);
    }
}

template <typename Policy, typename Problem, typename Functor>
__global__ void __launch_bounds__(Policy::THREADS, Policy::MIN_CTAS)
Kernel(const __grid_constant__ typename Problem::DataSlice ds,
       const __grid_constant__ PullArgs<typename Policy::VertexId, typename Policy::SizeT> args)
{
    __shared__ SweepStorage<Policy> s_sweep;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (blockIdx.x == 0 && threadIdx.x == 0) ClearFutureSlot(args.slots, args.queue_index);
    Sweep<Policy, Problem, Functor>(ds, args, s_sweep, slot_out,
                                    (long long)blockIdx.x * Policy::WARPS + util::WarpId(),
                                    (long long)gridDim.x * Policy::WARPS);
}

}  // namespace pull
}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
