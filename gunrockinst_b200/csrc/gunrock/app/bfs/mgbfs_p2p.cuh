// mgbfs_p2p.cuh -- peer-memory exchange of the 1D-partitioned BFS: frontier segments, discovery
// segments, level counters and the barriers between them travel as plain stores into the other
// devices' buffers over NVLink / NVSwitch (CUDA IPC mappings of each process's own
// allocations), issued by the kernels that produce the data -- no collective library call and
// no extra pass between "compute" and "exchange".
//
// No reference counterpart: the reference only declares the striping rule of a multi-GPU
// problem (gunrock/app/problem_base.cuh:186-210) and leaves the path as a TODO (:336-338).
// The NCCL exchange in gunrockinst_b200/partitioned.py is the baseline this replaces.
//
// Every device owns one Mailbox that all peers can write:
//     ctr[parity][sender]   the sender's 64-byte counter block of a level
//     flag[parity][sender]  epoch of that block (monotone, never reset)
//     bar[sender]           mid-level barrier count of the sender (monotone)
// A level ends with ExchangeCountersKernel: post own block + flag into every mailbox (own
// included), wait until every sender's flag has reached the epoch, copy all blocks to pinned
// host memory.  That is the barrier of the level as well: a device passes it only after every
// device has finished the level's kernels, i.e. after every remote store into its buffers.
// Blocks are double buffered by epoch parity because a fast device may post epoch e+1 while a
// slow one is still copying epoch e (it cannot get further ahead: e+1 completes only with the
// slow device's own e+1 post).
//
// Memory ordering: data stores of earlier kernels precede the flag store in stream order; the
// poster additionally issues __threadfence_system() before a release store of the flag, the
// waiter reads flags with acquire loads and fences before the data is consumed.
#pragma once
#include "../../oprtr/work_progress.cuh"
#include <cstdlib>
#include "../../util/device_intrinsics.cuh"

namespace gunrock {
namespace app {
namespace bfs {

constexpr int kMaxPeers = 8;

struct Mailbox {
    oprtr::LevelCounters ctr[2][kMaxPeers];
    unsigned long long   flag[2][kMaxPeers];
    unsigned long long   bar[kMaxPeers];
    unsigned long long   timed_out;         // set by a waiter that gave up
};

// what a kernel needs to reach the peers (kernel parameter; pointers are IPC mappings, own
// rank's entry is the local allocation)
struct MailboxTable {
    int       parts = 1, rank = 0;
    Mailbox  *mailbox[kMaxPeers]  = {};
};
struct PeerTable : MailboxTable {
    unsigned *map[2][kMaxPeers]   = {};     // the two global frontier maps of every device
    unsigned *recv[kMaxPeers]     = {};     // received discovery segments of every device
};

__device__ __forceinline__ void StoreReleaseSys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long LoadAcquireSys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long PeerTimerNs()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// spin until *flag >= want; gives up after `budget_ns` (a peer died or the ranks diverged) so
// that a broken run ends with an error instead of a hung device
__device__ __forceinline__ bool WaitAtLeast(const unsigned long long *flag, unsigned long long want,
                                            unsigned long long budget_ns)
{
    const unsigned long long t0 = PeerTimerNs();
    unsigned spins = 0;
    while (LoadAcquireSys(flag) < want) {
        if ((++spins & 1023u) == 0 && PeerTimerNs() - t0 > budget_ns) return false;
    }
    return true;
}

constexpr unsigned long long kWaitBudgetNs = 4000000000ull;      // 4 s

// <<<1, 32 * parts... one CTA of kMaxPeers warps>>>: warp p talks to peer p.
// Posts `mine` (epoch `epoch`) everywhere, waits for everybody's post, copies the blocks to
// `h_all` (pinned host) in rank order.
static __global__ void __launch_bounds__(32 * kMaxPeers)
ExchangeCountersKernel(MailboxTable pt, const oprtr::LevelCounters *mine, unsigned long long epoch,
                       oprtr::LevelCounters *h_all, unsigned long long *h_status,
                       unsigned long long budget_ns = kWaitBudgetNs)
{
    const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int par = (int)(epoch & 1);
    __shared__ int s_bad;
    if (threadIdx.x == 0) s_bad = 0;
    __syncthreads();
    if (p < pt.parts) {
        Mailbox *mb = pt.mailbox[p];
        const unsigned long long *src = reinterpret_cast<const unsigned long long *>(mine);
        unsigned long long *dst = reinterpret_cast<unsigned long long *>(&mb->ctr[par][pt.rank]);
        if (lane < 8) dst[lane] = src[lane];
        __threadfence_system();
        __syncwarp();
        if (lane == 0) StoreReleaseSys(&mb->flag[par][pt.rank], epoch);
        // ... and wait for peer p's post into MY mailbox
        Mailbox *own = pt.mailbox[pt.rank];
        bool ok = true;
        if (lane == 0) ok = WaitAtLeast(&own->flag[par][p], epoch, budget_ns);
        ok = __shfl_sync(0xffffffffu, ok, 0);
        if (!ok) { if (lane == 0) { s_bad = 1; own->timed_out = epoch; } }
        __threadfence_system();
        const unsigned long long *got = reinterpret_cast<const unsigned long long *>(&own->ctr[par][p]);
        if (lane < 8) reinterpret_cast<unsigned long long *>(h_all + p)[lane] =
            *reinterpret_cast<const volatile unsigned long long *>(got + lane);
    }
    __syncthreads();
    // a mid-level barrier (PeerBarrierKernel) that gave up earlier in this level left its mark in the own
    // mailbox: the level's data may be incomplete, so the exchange that ends the level reports it to the
    // host -- the mark is sticky, the context stays failed
    if (threadIdx.x == 0) {
        const bool late = *reinterpret_cast<const volatile unsigned long long *>(&pt.mailbox[pt.rank]->timed_out) != 0;
        __threadfence_system();
        *h_status = (s_bad || late) ? ~0ull : epoch;
    }
}

// barrier in the middle of a level (between "scatter my discoveries" and "absorb what I
// received"): bump my count in every mailbox, wait until every sender's count reached `count`
static __global__ void __launch_bounds__(32 * kMaxPeers)
PeerBarrierKernel(MailboxTable pt, unsigned long long count, unsigned long long budget_ns = kWaitBudgetNs)
{
    const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (p < pt.parts && lane == 0) {
        __threadfence_system();
        StoreReleaseSys(&pt.mailbox[p]->bar[pt.rank], count);
        Mailbox *own = pt.mailbox[pt.rank];
        if (!WaitAtLeast(&own->bar[p], count, budget_ns)) own->timed_out = count;
        __threadfence_system();
    }
}

// push level, exchange step: every 128-byte line of the local discovery map that holds a bit
// goes to its owner's receive buffer (slot `rank` of it) as one coalesced store, and is cleared
// locally for the next level.  Lines without discoveries cost a local read only, so a sparse
// level sends a few lines and a dense one at most the whole map.
static __global__ void ScatterNextKernel(PeerTable pt, unsigned *next, unsigned long long seg_words,
                                         unsigned long long *sent_words)
{
    const unsigned long long words = seg_words * (unsigned long long)pt.parts;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long mine = 0;
    for (unsigned long long w0 = (unsigned long long)blockIdx.x * blockDim.x; w0 < words; w0 += stride) {
        const unsigned long long w = w0 + threadIdx.x;
        const unsigned x = (w < words) ? next[w] : 0u;
        if (__ballot_sync(0xffffffffu, x != 0u) && w < words) {     // seg_words is a multiple of 32:
            const unsigned long long owner = w / seg_words;         // a warp's line has one owner
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

// first pull level after push levels: every device copies its own segment of frontier map
// `cur` into that map on every peer
static __global__ void ScatterSegmentKernel(PeerTable pt, int cur, unsigned long long seg_words)
{
    const unsigned *src = pt.map[cur][pt.rank] + (unsigned long long)pt.rank * seg_words;
    const unsigned long long vecs = seg_words >> 2;                  // seg_words % 32 == 0
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < vecs; i += stride) {
        const uint4 v = reinterpret_cast<const uint4 *>(src)[i];
        for (int p = 0; p < pt.parts; ++p) {
            if (p == pt.rank) continue;
            reinterpret_cast<uint4 *>(pt.map[cur][p] + (unsigned long long)pt.rank * seg_words)[i] = v;
        }
    }
}

// seeds the counter block the first exchange of a search carries: frontier = {src} on its
// owner, out_degree = deg(src)
template <typename SizeT>
__global__ void SeedCountersKernel(oprtr::LevelCounters *c, const SizeT *row_offsets, long long row)
{
    c->out_len = row >= 0 ? 1ull : 0ull;
    c->out_degree = row >= 0 ? (unsigned long long)(row_offsets[row + 1] - row_offsets[row]) : 0ull;
    c->edges = 0; c->scan_total = 0; c->aux0 = 0; c->aux1 = 0; c->overflow = 0; c->reserved = 0;
}

// labels <- -1, source label, source queue entry, source bit in the visited map and in frontier
// map 0 (own segment cleared first): the per-search fills of Problem::Reset without its host
// round trips.  Bitmap clears are cudaMemsetAsync calls issued by the caller.
template <typename VertexId, typename DataSlice>
__global__ void SeedSourceKernel(DataSlice ds, VertexId src, VertexId *queue, unsigned *frontier_map0)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    if (src < 0 || (long long)src >= ds.nodes) return;
    const unsigned long long pos = ds.BitPos(src);
    ds.d_visited_mask[pos >> 5] = 1u << (pos & 31);
    if (ds.part.Mine(src)) {
        ds.d_labels[ds.part.Local(src)] = 0;
        queue[0] = src;
        frontier_map0[pos >> 5] = 1u << (pos & 31);
    }
}


// Host side of the mailbox protocol for one device: owns the mailbox and the pinned landing
// area, counts epochs.  (MGBFSEnactor carries the same state inline; new partitioned
// primitives use this.)
struct PeerComm {
    MailboxTable table;
    util::DeviceArray<Mailbox> mailbox;
    util::PinnedArray<oprtr::LevelCounters> h_all;
    util::PinnedArray<unsigned long long> h_status;
    unsigned long long epoch = 0, bar_count = 0;
    unsigned long long budget_ns = kWaitBudgetNs;       // GRB_PEER_TIMEOUT_MS overrides (tests)
    bool ready = false;

    static unsigned long long BudgetFromEnv()
    {
        const char *e = getenv("GRB_PEER_TIMEOUT_MS");
        const long long ms = e ? atoll(e) : 0;
        return ms > 0 ? (unsigned long long)ms * 1000000ull : kWaitBudgetNs;
    }

    cudaError_t Setup(int parts, int rank, cudaStream_t s)
    {
        budget_ns = BudgetFromEnv();
        if (parts > kMaxPeers) return util::GRError(cudaErrorInvalidValue, "peer exchange supports up to 8 devices", __FILE__, __LINE__);
        GR_TRY(mailbox.Allocate(1));
        GR_TRY(mailbox.Zero(s));
        GR_TRY(h_all.Allocate(kMaxPeers));
        GR_TRY(h_status.Allocate(1));
        table = MailboxTable();
        table.parts = parts; table.rank = rank;
        table.mailbox[rank] = mailbox.ptr;
        epoch = bar_count = 0;
        ready = (parts == 1);
        return cudaStreamSynchronize(s);
    }
    // posts `mine`, waits for everybody's block of this epoch; h_all.ptr[0..parts) valid on return
    cudaError_t Exchange(const oprtr::LevelCounters *mine, cudaStream_t s)
    {
        ++epoch;
        *h_status.ptr = 0;
        ExchangeCountersKernel<<<1, 32 * kMaxPeers, 0, s>>>(table, mine, epoch, h_all.ptr, h_status.ptr, budget_ns);
        GR_TRY(cudaGetLastError());
        volatile unsigned long long *st = h_status.ptr;
        unsigned long long spins = 0;
        while (*st == 0) { if (++spins > (1ull << 33)) break; }
        if (*st == 0) GR_TRY(cudaStreamSynchronize(s));
        if (*h_status.ptr != epoch)
            return util::GRError(cudaErrorLaunchTimeout, "peer exchange timed out (a peer stopped or the ranks diverged)", __FILE__, __LINE__);
        return cudaSuccess;
    }
    cudaError_t Barrier(cudaStream_t s)
    {
        if (table.parts == 1) return cudaSuccess;
        ++bar_count;
        PeerBarrierKernel<<<1, 32 * kMaxPeers, 0, s>>>(table, bar_count, budget_ns);
        return cudaGetLastError();      // a barrier that gives up is reported by the Exchange that ends the level
    }
};

}  // namespace bfs
}  // namespace app
}  // namespace gunrock
