// noalloc_capi.cu -- grb_advance_lb_noalloc (include/gunrock_b200.h): the load-balanced advance with
// the probe functor of oprtr_capi.cu, for callers that cannot allocate device memory (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.
//
// The operator is the one grb_test_advance(mode = 1) runs -- lb::ScanKernel then lb::ExpandKernel, same
// policy, same functor -- but
//   * every array argument is a DEVICE pointer owned by the caller, the work is queued on the caller's
//     stream and nothing is copied to or from the host;
//   * the scratch of the LB mode (degree scan, row starts, look-back descriptors, tile split points,
//     scan ticket, the four rotating counter slots) lives in ONE file-scope __device__ array of
//     kScratchBytes (about 4.2 MiB) that every GPU which loads the library holds until the process
//     exits.  Its device address is looked up once per device with cudaGetSymbolAddress (below); the
//     array's name is only a host-side stand-in in host code and is never handed to a kernel;
//   * the entry point calls no cudaMalloc / cudaFree / cudaMallocHost and none of the library's own
//     allocating helpers (util::DeviceArray, WorkProgress::Setup, LbScratch::Allocate).
// Limits: GRB_NOALLOC_MAX_ITEMS input items, GRB_NOALLOC_MAX_EDGES neighbour entries per call; a longer
// input is rejected up front, a frontier that expands past the edge limit (the degree sum is only known
// on the device) raises the `overflow` counter, exactly like an output queue that is too small.
// Every call shares the array and finds the previous call's contents in it: what must start at zero
// (counter slots, ticket, descriptors) is cleared on the caller's stream before the kernels.  Two calls
// that overlap on the GPU (different streams without an ordering between them, or two host threads on one
// device) RACE on the scratch; callers serialise them.
#include <cstdio>
#include <mutex>
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/app/enactor_base.cuh"
#include "../gunrock/app/problem_base.cuh"

namespace {

constexpr unsigned long long kMaxItems = GRB_NOALLOC_MAX_ITEMS;
constexpr unsigned long long kMaxEdges = GRB_NOALLOC_MAX_EDGES;
constexpr int kScanTile  = 1024;        // Policy::SCAN_TILE, checked below
constexpr int kTileEdges = 1024;        // Policy::TILE_EDGES

constexpr size_t Align256(size_t x) { return (x + 255) & ~(size_t)255; }
// layout of the scratch array (byte offsets, each part on a 256-byte boundary)
constexpr size_t kOffSlots  = 0;
constexpr size_t kOffTicket = kOffSlots  + Align256(4 * sizeof(gunrock::oprtr::LevelCounters));
constexpr size_t kOffDesc   = kOffTicket + Align256(sizeof(unsigned));
constexpr size_t kDescWords = kMaxItems / kScanTile + 2;
constexpr size_t kOffPart   = kOffDesc   + Align256(kDescWords * sizeof(unsigned long long));
constexpr size_t kPartWords = kMaxEdges / kTileEdges + 2;
constexpr size_t kOffScan   = kOffPart   + Align256(kPartWords * sizeof(long long));
constexpr size_t kOffRow    = kOffScan   + Align256((kMaxItems + 1) * sizeof(unsigned long long));
constexpr size_t kScratchBytes = kOffRow + Align256((kMaxItems + 1) * sizeof(long long));    // row starts sized for 64-bit offsets
static_assert(kScratchBytes <= (size_t)GRB_NOALLOC_SCRATCH_BYTES, "the header states the size of the scratch array");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// device address of the array on each device, resolved on first use with that device current
constexpr int kMaxDevices = 64;
std::mutex     g_lock;
unsigned char *g_scratch_ptr[kMaxDevices] = {};

cudaError_t ScratchOnDevice(int device, unsigned char **out)
{
    if (device < 0 || device >= kMaxDevices)
        return gunrock::util::GRError(cudaErrorInvalidDevice, "grb_advance_lb_noalloc: device index beyond 63", __FILE__, __LINE__);
    std::lock_guard<std::mutex> guard(g_lock);
    if (!g_scratch_ptr[device]) {
        int current = -1;
        GR_TRY(cudaGetDevice(&current));
        if (current != device)
            return gunrock::util::GRError(cudaErrorInvalidDevice, "grb_advance_lb_noalloc: the graph's device is not current", __FILE__, __LINE__);
        void *p = nullptr;
        GR_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return gunrock::util::GRError(cudaErrorInvalidSymbol, "grb_advance_lb_noalloc: scratch symbol has no address", __FILE__, __LINE__);
        g_scratch_ptr[device] = static_cast<unsigned char *>(p);
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

// the probe problem / functor of oprtr_capi.cu (grb_test_advance), restated here so that the existing
// translation unit stays as it is
template <typename _SizeT>
struct ProbeProblem {
    typedef int    VertexId;
    typedef _SizeT SizeT;
    typedef int    Value;
    static const bool MARK_PREDECESSORS  = true;
    static const bool ENABLE_IDEMPOTENCE = false;
    struct DataSlice {
        int       *edge_hits;
        long long *src_sum;
        int       *dst_hits;
        long long  nodes, local_nodes;
        gunrock::graph::Partition part;
    };
};

template <typename Problem>
struct ProbeFunctor {
    typedef typename Problem::DataSlice DataSlice;
    typedef typename Problem::SizeT SizeT;
    static __device__ __forceinline__ bool CondEdge(int s, int d, const DataSlice *p, SizeT e = 0, SizeT e_in = 0)
    {
        atomicAdd(p->edge_hits + e, 1);
        atomicAdd((unsigned long long *)p->src_sum + d, (unsigned long long)(long long)s);
        return ((s + d) % 3) != 0;
    }
    static __device__ __forceinline__ void ApplyEdge(int s, int d, const DataSlice *p, SizeT e = 0, SizeT e_in = 0)
    {
        atomicAdd(p->dst_hits + d, 1);
    }
    static __device__ __forceinline__ bool CondFilter(int v, const DataSlice *p, int val = 0, SizeT nid = 0) { return (v % 2) == 0; }
    static __device__ __forceinline__ void ApplyFilter(int v, const DataSlice *p, int val = 0, SizeT nid = 0) { atomicAdd(p->dst_hits + v, 1); }
};

template <typename SizeT>
cudaError_t RunNoAlloc(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t s, int output,
                       const int32_t *in_queue, int64_t in_len, int32_t *out_queue, int64_t out_cap,
                       int32_t *edge_hits, int64_t *src_sum, int32_t *dst_hits, uint64_t *counters)
{
    using namespace gunrock;
    typedef ProbeProblem<SizeT> Problem;
    typedef ProbeFunctor<Problem> Functor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, false> Policy;     // grb_test_advance's policy
    static_assert(Policy::SCAN_TILE == kScanTile && Policy::TILE_EDGES == kTileEdges, "scratch layout follows the policy");
    static_assert(sizeof(SizeT) <= sizeof(long long), "row starts fit their part of the scratch");
    namespace lb = oprtr::advance::lb;

    unsigned char *base = nullptr;
    GR_TRY(ScratchOnDevice(g->device, &base));
    oprtr::advance::DeviceInfo dev;
    GR_TRY(dev.Query(g->device));

    // the previous call's counters, ticket and descriptors are still in the array: slots .. descriptors are contiguous
    GR_TRY(cudaMemsetAsync(base + kOffSlots, 0, kOffPart - kOffSlots, s));

    oprtr::advance::LbScratchView<SizeT> sv;
    sv.scanned    = reinterpret_cast<unsigned long long *>(base + kOffScan);
    sv.row_start  = reinterpret_cast<SizeT *>(base + kOffRow);
    sv.tile_desc  = reinterpret_cast<unsigned long long *>(base + kOffDesc);
    sv.part_start = reinterpret_cast<long long *>(base + kOffPart);
    sv.ticket     = reinterpret_cast<unsigned *>(base + kOffTicket);
    sv.part_cap   = kPartWords;
    sv.epoch      = 1;

    typename Problem::DataSlice ds;
    ds.edge_hits = edge_hits; ds.src_sum = reinterpret_cast<long long *>(src_sum); ds.dst_hits = dst_hits;
    ds.nodes = g->nodes; ds.local_nodes = g->local_nodes; ds.part = g->part;
    oprtr::advance::Args<int, SizeT> a;
    a.g = g->View();
    a.in_queue = in_queue;
    a.out_queue = output ? out_queue : nullptr;
    a.in_len = (unsigned long long)in_len; a.d_in_len = nullptr;
    a.out_cap = (unsigned long long)out_cap;
    a.slots = reinterpret_cast<oprtr::LevelCounters *>(base + kOffSlots);
    a.queue_index = 0; a.iteration = 7;
    a.type = oprtr::advance::V2V;
    a.output = output ? oprtr::advance::OUT_COMPACT : oprtr::advance::OUT_NONE;

    const unsigned long long scan_tiles = ((unsigned long long)in_len + Policy::SCAN_TILE - 1) / Policy::SCAN_TILE;
    lb::ScanKernel<Policy><<<(unsigned)(scan_tiles < 1 ? 1 : scan_tiles), Policy::SCAN_THREADS, 0, s>>>(a, sv);
    GR_TRY(cudaGetLastError());
    const unsigned long long resident = (unsigned long long)dev.sm_count * Policy::MIN_CTAS;
    unsigned long long hint = (unsigned long long)g->edges < kMaxEdges ? (unsigned long long)g->edges : kMaxEdges;
    const unsigned long long tiles = (hint + Policy::TILE_EDGES - 1) / Policy::TILE_EDGES;
    lb::ExpandKernel<Policy, Problem, Functor><<<(unsigned)(tiles < 1 ? 1 : (tiles < resident ? tiles : resident)), Policy::THREADS, 0, s>>>(ds, a, sv);
    GR_TRY(cudaGetLastError());
    // the slot the kernels accumulated into (queue_index + 1) -> the caller's eight counters
    if (counters)
        GR_TRY(cudaMemcpyAsync(counters, a.slots + 1, sizeof(oprtr::LevelCounters), cudaMemcpyDeviceToDevice, s));
    return cudaSuccess;
}

}  // namespace

extern "C" int grb_advance_lb_noalloc(grb_graph_t g, int output, const int32_t *in_queue_dev, int64_t in_len,
                                      int32_t *out_queue_dev, int64_t out_cap,
                                      int32_t *edge_hits_dev, int64_t *src_sum_dev, int32_t *dst_hits_dev,
                                      uint64_t *counters_dev, void *cuda_stream)
{
    using gunrock::util::GRError;
    if (!g || in_len < 0 || out_cap < 0 || !in_queue_dev || !edge_hits_dev || !src_sum_dev || !dst_hits_dev ||
        (output != 0 && output != 1) || (output == 1 && !out_queue_dev))
        return (int)GRError(cudaErrorInvalidValue, "grb_advance_lb_noalloc: bad argument", __FILE__, __LINE__);
    if ((unsigned long long)in_len > kMaxItems) {
        fprintf(stderr, "grb_advance_lb_noalloc: in_len %lld exceeds GRB_NOALLOC_MAX_ITEMS (%llu): the static scratch of this "
                        "entry point cannot hold the frontier; use the allocating entry points\n", (long long)in_len, kMaxItems);
        return (int)cudaErrorInvalidValue;
    }
    if (g->g32 ? g->g32->part.log_parts != 0 : g->g64->part.log_parts != 0)
        return (int)GRError(cudaErrorInvalidValue, "grb_advance_lb_noalloc: whole graphs only (not a slice of a partition)", __FILE__, __LINE__);
    GRB_TRY(cudaSetDevice(g->device()));
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    return (int)(g->g32 ? RunNoAlloc<int>(g->g32, s, output, in_queue_dev, in_len, out_queue_dev, out_cap,
                                          edge_hits_dev, src_sum_dev, dst_hits_dev, counters_dev)
                        : RunNoAlloc<long long>(g->g64, s, output, in_queue_dev, in_len, out_queue_dev, out_cap,
                                                edge_hits_dev, src_sum_dev, dst_hits_dev, counters_dev));
}
