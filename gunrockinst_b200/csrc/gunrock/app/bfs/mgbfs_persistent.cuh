// mgbfs_persistent.cuh -- a whole search of the 1D-partitioned, direction-optimising BFS as ONE cooperative
// kernel per device.
//
// The host loop of MGBFSEnactor::RunPeer drives a level with ~5 launches, a single-CTA counter-exchange kernel
// and a poll of a pinned status word (~40 launches per search and device; a level that expands 109 vertices
// cost 70 us, profiles/r01w_peer_n8_s27.json).  Here the level loop runs on the devices, the way
// bfs_persistent.cuh does it for one device:
//
//   * every device runs this kernel on 148 x k co-resident CTAs; the phases of a level are the SAME operator
//     bodies the stand-alone kernels wrap (advance::lb::ScanTiles / ExpandTiles, twc::ExpandItems,
//     pull::Sweep with its stores into the peers' frontier maps, the scatter / absorb / merge bodies of
//     mgbfs_enactor.cuh and mgbfs_p2p.cuh), separated by grid-wide barriers;
//   * the barriers BETWEEN devices and the counter exchange that ends a level are the mailbox protocol of
//     mgbfs_p2p.cuh, executed by CTA 0 (warp p talks to peer p) between two grid-wide barriers: every thread
//     fences its remote stores (fence.sys), the grid meets, CTA 0 posts flag / counters with release stores
//     into every mailbox and spins (acquire loads) until every peer has posted, the grid meets again and every
//     thread reads all devices' 64-byte counter blocks from device memory;
//   * so every thread of every device holds the same global figures (frontier length, its degree sum) and
//     takes the same decisions -- termination, Beamer's push <-> pull switch, LB or TWC -- with no host in the
//     loop.  The host launches once per search and reads one summary.
//
// A waiter gives up after the mailbox budget (an error, not a hang): the flag is sticky, every device sees a
// failed exchange in the same epoch or times out itself, and the kernel returns with summary.error set.
//
// No reference counterpart (the reference's multi-GPU path is a TODO, gunrock/app/problem_base.cuh:336-338).
#pragma once
#include <cooperative_groups.h>
#include "bfs_persistent.cuh"
#include "mgbfs_p2p.cuh"

namespace gunrock {
namespace app {
namespace bfs {

struct MgLevelRecord {
    int level, kind;                                // kind: 0 push, 1 pull
    unsigned long long frontier, discovered, edges, max_edges, candidates;
    unsigned long long ns_compute, ns_exchange;     // globaltimer spans seen by the first thread of the device
    unsigned long long sent_words;                  // push: words this device had stored into peers when the level ended
};

struct MgPersistentSummary {
    unsigned long long levels, push_levels, pull_levels;
    unsigned long long edges_visited, edges_inspected, nodes_visited;
    unsigned long long mailbox_epoch, barrier_count;    // protocol counters after the search (the host carries them on)
    unsigned long long lb_epoch_end, in_sel;
    unsigned long long error;                           // 1: a peer did not answer in time, 2: queue overflow
    unsigned long long ns_kernel, grid_syncs;
    unsigned long long gather_bytes;                    // frontier-map bytes this device stored into peers (pull levels)
    // TAIL instance: the loop state when the kernel returned (F == 0: the search is over; handoff != 0: the host loop goes on)
    unsigned long long handoff, F, frontier_edges, local_q, local_hint, prev_len, unexplored, level, cur, queue_valid, gathered;
};

// what the host loop hands to the TAIL instance (the state after its last counter exchange)
struct MgTailState {
    unsigned long long F = 0, frontier_edges = 0, local_q = 0, local_hint = 0, prev_len = 0, unexplored = 0, limit = 0;
    int level = 0, cur = 0, queue_valid = 1, gathered = 0;
};

template <typename VertexId, typename SizeT>
struct MgPersistentParams {
    graph::CsrView<VertexId, SizeT> g, ginv;
    const SizeT *fwd_row_offsets;
    VertexId    *queues[2];
    unsigned long long queue_cap[2];
    unsigned    *frontier_map[2];
    unsigned    *visited, *next_map, *recv_map;
    const unsigned *candidates;
    const VertexId *head;
    oprtr::advance::LbScratchView<SizeT> scratch;       // .epoch = last epoch used
    oprtr::LevelCounters *slots;
    PeerTable    peers;
    oprtr::LevelCounters *all;                          // [kMaxPeers] device memory: every device's block of the level
    unsigned long long *status;                         // device memory: != 0 after a failed exchange / barrier
    unsigned long long *sent_words;
    MgLevelRecord *records;
    int          max_records;
    MgPersistentSummary *summary;
    unsigned long long mailbox_epoch, barrier_count, budget_ns, global_edges, seg_words, desc_words;
    int          dopt, in_sel;
    double       alpha, beta;
    unsigned long long lb_min_edges;
    MgTailState  tail;
};

// ---- bodies shared with the stand-alone kernels of mgbfs_enactor.cuh / mgbfs_p2p.cuh (grid-stride over `nctas` CTAs) ----

// visited |= map over the whole global bitmap
__device__ __forceinline__ void MergeVisitedBody(unsigned *visited, const unsigned *map, unsigned long long words,
                                                 unsigned cta, unsigned nctas)
{
    const unsigned long long stride = (unsigned long long)nctas * blockDim.x;
    for (unsigned long long w = (unsigned long long)cta * blockDim.x + threadIdx.x; w < words; w += stride) {
        const unsigned m = map[w];
        if (m) visited[w] |= m;
    }
}

// every 128-byte line of the local discovery map that holds a bit goes to slot `rank` of its owner's receive buffer
__device__ __forceinline__ void ScatterNextBody(const PeerTable &pt, unsigned *next, unsigned long long seg_words,
                                                unsigned long long *sent_words, unsigned cta, unsigned nctas)
{
    const unsigned long long words = seg_words * (unsigned long long)pt.parts;
    const unsigned long long stride = (unsigned long long)nctas * blockDim.x;
    unsigned long long mine = 0;
    for (unsigned long long w0 = (unsigned long long)cta * blockDim.x; w0 < words; w0 += stride) {
        const unsigned long long w = w0 + threadIdx.x;
        const unsigned x = (w < words) ? next[w] : 0u;
        if (__ballot_sync(0xffffffffu, x != 0u) && w < words) {
            const unsigned long long owner = w / seg_words;
            const unsigned long long off = w - owner * seg_words;
            pt.recv[owner][(unsigned long long)pt.rank * seg_words + off] = x;
            if (x) next[w] = 0;
            if ((int)owner != pt.rank) ++mine;
        }
    }
    if (sent_words) {
        for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
        if ((threadIdx.x & 31) == 0 && mine) atomicAdd(sent_words, mine);
    }
}

// own segment of frontier map `cur` -> that map on every peer
__device__ __forceinline__ void ScatterSegmentBody(const PeerTable &pt, int cur, unsigned long long seg_words,
                                                   unsigned cta, unsigned nctas)
{
    const unsigned *src = pt.map[cur][pt.rank] + (unsigned long long)pt.rank * seg_words;
    const unsigned long long vecs = seg_words >> 2;
    const unsigned long long stride = (unsigned long long)nctas * blockDim.x;
    for (unsigned long long i = (unsigned long long)cta * blockDim.x + threadIdx.x; i < vecs; i += stride) {
        const uint4 v = reinterpret_cast<const uint4 *>(src)[i];
        for (int p = 0; p < pt.parts; ++p) {
            if (p == pt.rank) continue;
            reinterpret_cast<uint4 *>(pt.map[cur][p] + (unsigned long long)pt.rank * seg_words)[i] = v;
        }
    }
}

template <int THREADS>
struct AbsorbStorage {
    unsigned           warp[THREADS / 32 + 1];
    unsigned long long red[THREADS / 32 + 1];
    unsigned long long base;
};

// OR the received copies of this device's segment, keep what is new: labels, own frontier segment, local queue, counters
template <typename VertexId, typename SizeT, typename DataSlice, int THREADS>
__device__ __forceinline__ void AbsorbDenseBody(unsigned *recv, int parts, bool consume, unsigned *visited, unsigned *frontier_out,
                                                VertexId *queue, unsigned long long cap, const SizeT *row_offsets,
                                                oprtr::LevelCounters *slot, const DataSlice &ds, AbsorbStorage<THREADS> &sm,
                                                unsigned cta, unsigned nctas)
{
    const long long words = (ds.local_nodes + 31) >> 5;
    const long long seg = ((long long)ds.part.rank * ds.local_nodes) >> 5;
    const long long rounds = (words + (long long)nctas * THREADS - 1) / ((long long)nctas * THREADS);
    unsigned long long my_degree = 0;
    for (long long r = 0; r < rounds; ++r) {
        const long long w = (r * nctas + cta) * THREADS + threadIdx.x;
        unsigned bits = 0;
        if (w < words) {
            unsigned acc = 0;
            for (int p = 0; p < parts; ++p) {
                const unsigned x = util::LoadStream(recv + (long long)p * words + w);
                if (consume && x) recv[(long long)p * words + w] = 0;
                acc |= x;
            }
            const unsigned vis = visited[seg + w];
            bits = acc & ~vis;
            frontier_out[seg + w] = bits;
            if (bits) visited[seg + w] = vis | bits;
        }
        unsigned total;
        unsigned excl = util::BlockExclusiveSum<THREADS>((unsigned)__popc(bits), sm.warp, total);
        if (total == 0) continue;                       // uniform
        if (threadIdx.x == 0) sm.base = atomicAdd(&slot->out_len, (unsigned long long)total);
        __syncthreads();
        unsigned long long at = sm.base + excl;
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            const long long row = (w << 5) + b;
            util::StoreCG(ds.d_labels + row, (VertexId)(ds.level + 1));
            my_degree += (unsigned long long)(row_offsets[row + 1] - row_offsets[row]);
            if (at < cap) queue[at] = (VertexId)ds.part.Global(row); else slot->overflow = 1;
            ++at;
        }
        __syncthreads();
    }
    oprtr::advance::BlockAccumulate<THREADS>(my_degree, sm.red, &slot->out_degree);
}

// ---- the mailbox protocol from inside a kernel: the whole CTA calls these (>= 32 * parts threads), warp p talks to peer p ----

// post `mine` (epoch `epoch`) into every mailbox, wait for everybody's post, copy the blocks to `all` (device memory) in
// rank order; clears `mine` for the next level; *status != 0 afterwards if somebody did not answer in time
__device__ __forceinline__ void ExchangeCountersBody(const MailboxTable &pt, oprtr::LevelCounters *mine, unsigned long long epoch,
                                                     oprtr::LevelCounters *all, unsigned long long *status, unsigned long long budget_ns)
{
    const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int par = (int)(epoch & 1);
    unsigned long long word = 0;
    if (lane < 8) word = *reinterpret_cast<volatile unsigned long long *>(reinterpret_cast<unsigned long long *>(mine) + lane);
    __syncthreads();                                    // everybody holds the block: it can be cleared for the next level
    if (threadIdx.x < 8) reinterpret_cast<unsigned long long *>(mine)[threadIdx.x] = 0;
    if (p < pt.parts) {
        Mailbox *mb = pt.mailbox[p];
        unsigned long long *dst = reinterpret_cast<unsigned long long *>(&mb->ctr[par][pt.rank]);
        if (lane < 8) dst[lane] = word;
        __threadfence_system();
        __syncwarp();
        if (lane == 0) StoreReleaseSys(&mb->flag[par][pt.rank], epoch);
        Mailbox *own = pt.mailbox[pt.rank];
        bool ok = true;
        if (lane == 0) ok = WaitAtLeast(&own->flag[par][p], epoch, budget_ns);
        ok = __shfl_sync(0xffffffffu, ok, 0);
        if (!ok && lane == 0) { own->timed_out = epoch; *status = 1; }
        __threadfence_system();
        const unsigned long long *got = reinterpret_cast<const unsigned long long *>(&own->ctr[par][p]);
        if (lane < 8) reinterpret_cast<unsigned long long *>(all + p)[lane] =
            *reinterpret_cast<const volatile unsigned long long *>(got + lane);
    }
    __syncthreads();
    if (threadIdx.x == 0 && *reinterpret_cast<const volatile unsigned long long *>(&pt.mailbox[pt.rank]->timed_out) != 0) *status = 1;
    __threadfence();
}

// barrier in the middle of a level: bump my count in every mailbox, wait until every sender's count reached `count`
__device__ __forceinline__ void PeerBarrierBody(const MailboxTable &pt, unsigned long long count, unsigned long long *status,
                                                unsigned long long budget_ns)
{
    const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (p < pt.parts && lane == 0) {
        __threadfence_system();
        StoreReleaseSys(&pt.mailbox[p]->bar[pt.rank], count);
        Mailbox *own = pt.mailbox[pt.rank];
        if (!WaitAtLeast(&own->bar[p], count, budget_ns)) { own->timed_out = count; *status = 1; }
        __threadfence_system();
    }
    __syncthreads();
}

template <typename PushPolicy, typename PullPolicy>
union MgPersistentStorage {
    struct {
        oprtr::advance::lb::ExpandStorage<PushPolicy, false> expand;
        oprtr::advance::lb::SmallScanStorage<PushPolicy>     small;
    } push;
    oprtr::advance::lb::ScanStorage<PushPolicy::THREADS>   scan;
    oprtr::advance::twc::Storage<PushPolicy, false>        twc;
    oprtr::advance::pull::SweepStorage<PullPolicy>         sweep;
    struct { unsigned warp[PushPolicy::THREADS / 32 + 1]; unsigned long long base; } b2q;
    AbsorbStorage<PushPolicy::THREADS>                     absorb;
};

#ifndef GRB_MG_PERSISTENT_CTAS
#define GRB_MG_PERSISTENT_CTAS 4
#endif
constexpr int kMgPersistentCtasPerSm = GRB_MG_PERSISTENT_CTAS;

// TAIL = true is the instance the host loop uses for the END of a search: once the global frontier is short and shrinking
// (a few thousand vertices per level, where a level is two barriers between the devices and nothing else) the remaining
// push levels run here, top-down with the TWC expansion only -- no sweep, no LB scan in the kernel, so it is small -- until
// the frontier is empty, or until the direction rule asks for a pull level or the frontier grows past P.tail.limit again
// (then the loop state goes back to the host).
template <typename PushPolicy, typename PullPolicy, typename Problem, typename PushFunctor, typename PullFunctor, bool TAIL = false>
__global__ void __launch_bounds__(PushPolicy::THREADS, kMgPersistentCtasPerSm)
MgPersistentKernel(const __grid_constant__ typename Problem::DataSlice ds0,
                   const __grid_constant__ MgPersistentParams<typename Problem::VertexId, typename Problem::SizeT> P)
{
    namespace cg  = cooperative_groups;
    namespace adv = oprtr::advance;
    typedef typename Problem::VertexId  VertexId;
    typedef typename Problem::SizeT     SizeT;
    typedef typename Problem::DataSlice DataSlice;
    constexpr int THREADS = PushPolicy::THREADS;
    static_assert(PushPolicy::THREADS == PullPolicy::THREADS && THREADS >= 32 * kMaxPeers, "one CTA shape; a warp per peer");

    __shared__ MgPersistentStorage<PushPolicy, PullPolicy> sm;
    __shared__ unsigned long long s_level[8];
    cg::grid_group grid = cg::this_grid();
    const unsigned cta = blockIdx.x, num_ctas = gridDim.x;
    const bool leader = cta == 0 && threadIdx.x == 0;
    const unsigned long long gtid = (unsigned long long)cta * THREADS + threadIdx.x;
    const unsigned long long gthreads = (unsigned long long)num_ctas * THREADS;
    const int parts = P.peers.parts, rank = P.peers.rank;
    const unsigned long long n = (unsigned long long)P.g.nodes;
    const unsigned long long t_start = leader ? GlobalTimerNs() : 0ull;

    // ---- loop state: identical in every thread of every device --------------------------------------------
    DataSlice ds = ds0;
    oprtr::LevelCounters *slot = P.slots + 1;          // every phase of a level accumulates here (queue_index 0)
    unsigned long long epoch = P.mailbox_epoch, bar = P.barrier_count;
    unsigned lb_epoch = P.scratch.epoch;
    unsigned long long F = 0, frontier_edges = 0, local_q = 0, local_hint = 0, prev_len = 0, unexplored = 0;
    bool pull = false, gathered = false, queue_valid = true, first = true;
    int cur = 0, in_sel = P.in_sel;
    VertexId level = 0;
    MgPersistentSummary sum;
    sum.levels = sum.push_levels = sum.pull_levels = 0;
    sum.edges_visited = sum.edges_inspected = sum.nodes_visited = 0;
    sum.error = sum.grid_syncs = sum.gather_bytes = 0;
    sum.handoff = 0;
    unsigned long long t_level = 0, ns_exchange = 0;    // leader only
    unsigned long long lvl_frontier = 0;
    int lvl_kind = 0;
    bool entering = TAIL;                               // TAIL: the host's last exchange stands for this kernel's first
    if constexpr (TAIL) {
        F = P.tail.F; frontier_edges = P.tail.frontier_edges; local_q = P.tail.local_q; local_hint = P.tail.local_hint;
        prev_len = P.tail.prev_len; unexplored = P.tail.unexplored;
        level = (VertexId)P.tail.level; cur = P.tail.cur; queue_valid = P.tail.queue_valid != 0; gathered = P.tail.gathered != 0;
        first = false;
        // the host loop's counter block of the level before is still in the slot (its kernels clear it when a level starts)
        if (gtid < 8) reinterpret_cast<unsigned long long *>(slot)[gtid] = 0;
        grid.sync(); ++sum.grid_syncs;
    }

    for (;;) {
      const bool decided = entering;                    // TAIL: the host has taken the direction decision of the first level
      if (!entering) {
        // ---- the level ends (the search begins) with the counter exchange: the barrier between the devices --------
        const unsigned long long t_x0 = leader ? GlobalTimerNs() : 0ull;
        if (parts > 1) __threadfence_system();          // this thread's stores into peer memory, before the flag goes out
        grid.sync(); ++sum.grid_syncs;
        ++epoch;
        if (cta == 0) ExchangeCountersBody(P.peers, slot, epoch, P.all, P.status, P.budget_ns);
        grid.sync(); ++sum.grid_syncs;
        if (threadIdx.x == 0) {
            unsigned long long f = 0, fe = 0, ins = 0, max_ins = 0, ovf = 0, cands = 0;
            for (int p = 0; p < parts; ++p) {
                const oprtr::LevelCounters *c = P.all + p;
                f += util::LoadCG(&c->out_len); fe += util::LoadCG(&c->out_degree);
                const unsigned long long e = util::LoadCG(&c->edges);
                ins += e; max_ins = e > max_ins ? e : max_ins;
                ovf += util::LoadCG(&c->overflow); cands += util::LoadCG(&c->aux0);
            }
            s_level[0] = f; s_level[1] = fe; s_level[2] = ins; s_level[3] = ovf | (util::LoadCG(P.status) ? 4ull : 0ull);
            s_level[4] = util::LoadCG(&P.all[rank].out_len); s_level[5] = util::LoadCG(&P.all[rank].out_degree);
            s_level[6] = max_ins; s_level[7] = cands;
        }
        __syncthreads();
        const unsigned long long f = s_level[0], fe = s_level[1], ins = s_level[2], bad = s_level[3];
        const unsigned long long lq = s_level[4], lh = s_level[5], max_ins = s_level[6], cands = s_level[7];
        __syncthreads();
        if (leader) ns_exchange += GlobalTimerNs() - t_x0;
        if (bad) { sum.error = (bad & 4ull) ? 1 : 2; break; }
        if (first) {
            first = false;
            unexplored = P.global_edges > fe ? P.global_edges - fe : 0;
        } else {
            // bookkeeping of the level that has just ended
            if (leader && (int)sum.levels < P.max_records) {
                MgLevelRecord r;
                r.level = (int)level; r.kind = lvl_kind; r.frontier = lvl_frontier; r.discovered = f; r.edges = ins;
                r.max_edges = max_ins; r.candidates = lvl_kind ? cands : 0;
                const unsigned long long span = GlobalTimerNs() - t_level;
                r.ns_exchange = ns_exchange; r.ns_compute = span > ns_exchange ? span - ns_exchange : 0;
                r.sent_words = util::LoadCG(P.sent_words);
                P.records[sum.levels] = r;
            }
            ++sum.levels;
            sum.edges_inspected += ins;
            prev_len = F;
            unexplored = unexplored > fe ? unexplored - fe : 0;
            ++level;
        }
        F = f; frontier_edges = fe; local_q = lq; local_hint = lh;
        sum.edges_visited += fe;
        sum.nodes_visited += f;
        if (F == 0) break;
      }
        entering = false;
        if (leader) { t_level = GlobalTimerNs(); ns_exchange = 0; }
        lvl_frontier = F;

        // ---- direction (Beamer's rule on the global figures, as MGBFSEnactor::RunPeer) ---------------------------
        if constexpr (TAIL) {
            if (!decided && ((P.dopt && (double)frontier_edges > (double)unexplored / P.alpha) || F >= P.tail.limit)) { sum.handoff = 1; break; }
        } else if (P.dopt) {
            if (!pull && (double)frontier_edges > (double)unexplored / P.alpha) pull = true;
            else if (pull && F < prev_len && (double)F < (double)n / P.beta) pull = false;
        }
        ds.level = level;
        lvl_kind = pull ? 1 : 0;
        if constexpr (!TAIL) if (pull) {
            if (!gathered && parts > 1) {               // first pull level after push levels: own segment -> every peer
                const unsigned long long t0 = leader ? GlobalTimerNs() : 0ull;
                ScatterSegmentBody(P.peers, cur, P.seg_words, cta, num_ctas);
                __threadfence_system();
                grid.sync(); ++sum.grid_syncs;
                ++bar;
                if (cta == 0) PeerBarrierBody(P.peers, bar, P.status, P.budget_ns);
                grid.sync(); ++sum.grid_syncs;
                sum.gather_bytes += 4ull * P.seg_words * (unsigned long long)(parts - 1);
                if (leader) ns_exchange += GlobalTimerNs() - t0;
            }
            MergeVisitedBody(P.visited, P.frontier_map[cur], (n + 31) >> 5, cta, num_ctas);
            grid.sync(); ++sum.grid_syncs;
            adv::pull::PullArgs<VertexId, SizeT> pa;
            pa.g = P.ginv;
            pa.frontier_in  = P.frontier_map[cur];
            pa.frontier_out = P.frontier_map[cur ^ 1];
            pa.visited = P.visited;
            pa.candidates = P.candidates;
            pa.head = P.head;
            pa.tickets = nullptr;
            pa.fwd_row_offsets = P.fwd_row_offsets;
            pa.out_queue = nullptr; pa.out_cap = 0;
            pa.peers = 0;
            for (int p = 0; p < parts; ++p)
                if (p != rank) pa.peer_out[pa.peers++] = P.peers.map[cur ^ 1][p];
            pa.slots = P.slots; pa.queue_index = 0; pa.iteration = level;
            SweepCall<PullPolicy, Problem, PullFunctor>(ds, pa, sm.sweep, slot, (long long)cta * PullPolicy::WARPS + util::WarpId(),
                                                        (long long)num_ctas * PullPolicy::WARPS);
            sum.gather_bytes += 4ull * P.seg_words * (unsigned long long)(parts - 1);
            cur ^= 1;
            queue_valid = false;
            gathered = true;
            ++sum.pull_levels;
            continue;
        }
        {
            if (!queue_valid) {                         // own segment of map `cur` -> local queue (length known: local_q)
                if (leader) P.slots[3].out_len = 0;
                grid.sync(); ++sum.grid_syncs;
                BitmapToQueue<THREADS>(P.frontier_map[cur], P.queues[in_sel], P.queue_cap[in_sel], &P.slots[3].out_len,
                                       &slot->overflow, ds, sm.b2q, cta, num_ctas);
                grid.sync(); ++sum.grid_syncs;
                queue_valid = true;
            }
            if (local_q > 0) {                          // (uniform on this device; the other devices may skip or not)
                adv::Args<VertexId, SizeT> a;
                a.g = P.g;
                a.in_queue = P.queues[in_sel]; a.out_queue = nullptr;
                a.in_len = local_q; a.d_in_len = nullptr; a.out_cap = 0;
                a.slots = P.slots; a.queue_index = 0; a.iteration = level;
                a.output = adv::OUT_NONE;
                if (!TAIL && local_hint >= 4096) {
                  if constexpr (!TAIL) {
                    if (lb_epoch + 1 >= adv::lb::kEpochLimit) {
                        for (unsigned long long w = gtid; w < P.desc_words; w += gthreads) P.scratch.tile_desc[w] = 0;
                        grid.sync(); ++sum.grid_syncs;
                        lb_epoch = 0;
                    }
                    ++lb_epoch;
                    adv::LbScratchView<SizeT> sv = P.scratch;
                    sv.epoch = lb_epoch;
                    adv::lb::ScanTiles<PushPolicy, THREADS, 4>(a, sv, sm.scan, local_q, slot, true);
                    grid.sync(); ++sum.grid_syncs;
                    if (leader) *sv.ticket = 0;
                    adv::lb::ExpandTiles<PushPolicy, Problem, PushFunctor, false>(ds, a, sv, sm.push.expand, local_q, slot, cta, num_ctas);
                  }
                } else {
                    adv::twc::ExpandItems<PushPolicy, Problem, PushFunctor>(ds, a, sm.twc, local_q, slot, cta, num_ctas);
                }
            }
            grid.sync(); ++sum.grid_syncs;
            const unsigned long long t0 = leader ? GlobalTimerNs() : 0ull;
            ScatterNextBody(P.peers, P.next_map, P.seg_words, P.sent_words, cta, num_ctas);
            __threadfence_system();
            grid.sync(); ++sum.grid_syncs;
            ++bar;
            if (parts > 1 && cta == 0) PeerBarrierBody(P.peers, bar, P.status, P.budget_ns);
            grid.sync(); ++sum.grid_syncs;
            if (leader) ns_exchange += GlobalTimerNs() - t0;
            AbsorbDenseBody<VertexId, SizeT, DataSlice, THREADS>(P.recv_map, parts, true, P.visited, P.frontier_map[cur ^ 1],
                                                                 P.queues[in_sel ^ 1], P.queue_cap[in_sel ^ 1], P.g.row_offsets,
                                                                 slot, ds, sm.absorb, cta, num_ctas);
            in_sel ^= 1;
            cur ^= 1;
            gathered = false;
            ++sum.push_levels;
        }
    }
    if (leader) {
        if (!TAIL) sum.gather_bytes += 4ull * util::LoadCG(P.sent_words);     // + the discovery lines of the push levels = all peer stores
        sum.F = F; sum.frontier_edges = frontier_edges; sum.local_q = local_q; sum.local_hint = local_hint;
        sum.prev_len = prev_len; sum.unexplored = unexplored; sum.level = (unsigned long long)level; sum.cur = (unsigned long long)cur;
        sum.queue_valid = queue_valid ? 1 : 0; sum.gathered = gathered ? 1 : 0;
        sum.mailbox_epoch = epoch; sum.barrier_count = bar; sum.lb_epoch_end = lb_epoch; sum.in_sel = (unsigned long long)in_sel;
        sum.ns_kernel = GlobalTimerNs() - t_start;
        *P.summary = sum;
    }
}

}  // namespace bfs
}  // namespace app
}  // namespace gunrock
