// oprtr_capi.cu -- the two frontier operators, callable on their own through the C ABI
// (grb_test_advance / grb_test_filter, include/gunrock_b200.h) so that parity tests can check
// advance::LaunchKernel (reference: gunrock/oprtr/advance/kernel.cuh:101-129) and
// filter::Kernel (reference: gunrock/oprtr/filter/kernel.cuh:738-757) against a plain
// restatement of the operator contract, independent of any primitive.
//
// Probe functor (every side effect is an atomic add, so the result does not depend on the
// schedule):
//   CondEdge(s, d, e)   edge_hits[e] += 1; src_sum[d] += s; return (s + d) % 3 != 0
//   ApplyEdge(s, d, e)  dst_hits[d] += 1
//   CondFilter(v)       return v % 2 == 0          ApplyFilter(v)  dst_hits[v] += 1
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/app/enactor_base.cuh"
#include "../gunrock/app/problem_base.cuh"
#include "../gunrock/oprtr/row_bins.cuh"

namespace {

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
    static __device__ __forceinline__ bool CondFilter(int v, const DataSlice *p, int val = 0, SizeT nid = 0)
    {
        return (v % 2) == 0;
    }
    static __device__ __forceinline__ void ApplyFilter(int v, const DataSlice *p, int val = 0, SizeT nid = 0)
    {
        atomicAdd(p->dst_hits + v, 1);
    }
};

template <typename SizeT>
cudaError_t RunAdvance(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t s, int mode, int output,
                       const int32_t *in_queue, int64_t in_len, int32_t *out_queue, int64_t out_cap, int64_t *out_len,
                       int32_t *edge_hits, int64_t *src_sum, int32_t *dst_hits, uint64_t *edges_inspected)
{
    using namespace gunrock;
    typedef ProbeProblem<SizeT> Problem;
    typedef ProbeFunctor<Problem> Functor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, false> Policy;
    const size_t n = (size_t)g->nodes, m = (size_t)g->edges;
    util::DeviceArray<int> d_in, d_out, d_eh, d_dh;
    util::DeviceArray<long long> d_ss;
    oprtr::WorkProgress wp;
    oprtr::advance::LbScratch<SizeT> scratch;
    oprtr::advance::DeviceInfo dev;
    GR_TRY(dev.Query(g->device));
    GR_TRY(wp.Setup(s));
    if (in_queue) GR_TRY(d_in.FromHost(in_queue, (size_t)in_len, s));
    GR_TRY(d_out.Allocate((size_t)out_cap + 1));
    GR_TRY(cudaMemsetAsync(d_out.ptr, 0x7f, sizeof(int) * ((size_t)out_cap + 1), s));
    GR_TRY(d_eh.Allocate(m + 1)); GR_TRY(d_eh.Zero(s));
    GR_TRY(d_dh.Allocate(n + 1)); GR_TRY(d_dh.Zero(s));
    GR_TRY(d_ss.Allocate(n + 1)); GR_TRY(d_ss.Zero(s));
    const size_t items = (size_t)in_len + 1;
    GR_TRY(scratch.Allocate(items, m + 1, Policy::SCAN_TILE, Policy::TILE_EDGES, s));

    typename Problem::DataSlice ds;
    ds.edge_hits = d_eh.ptr; ds.src_sum = d_ss.ptr; ds.dst_hits = d_dh.ptr;
    ds.nodes = g->nodes; ds.local_nodes = g->local_nodes; ds.part = g->part;
    oprtr::advance::Args<int, SizeT> a;
    a.g = g->View();
    a.in_queue = in_queue ? d_in.ptr : nullptr;
    a.out_queue = output ? d_out.ptr : nullptr;
    a.in_len = (unsigned long long)in_len; a.d_in_len = nullptr;
    a.out_cap = (unsigned long long)out_cap;
    a.slots = wp.d_slots.ptr; a.queue_index = 0; a.iteration = 7;
    // bits 4-5 of `mode`: what the frontiers hold (0 V2V, 1 V2E, 2 E2V, 3 E2E)
    a.type = (oprtr::advance::TYPE)((mode >> 4) & 3);
    mode &= 15;
    a.output = output == 0 ? oprtr::advance::OUT_NONE : (output == 1 ? oprtr::advance::OUT_COMPACT : oprtr::advance::OUT_TOMBSTONE);
    if (mode == 0 && a.output == oprtr::advance::OUT_TOMBSTONE)
        return util::GRError(cudaErrorInvalidValue, "tombstone output exists in LB mode only", __FILE__, __LINE__);
    GR_TRY((oprtr::advance::LaunchKernel<Policy, Problem, Functor>(
        mode == 0 ? oprtr::advance::TWC_FORWARD : oprtr::advance::LB, dev, ds, a, &scratch, (unsigned long long)m, s)));
    oprtr::LevelCounters c;
    GR_TRY(wp.Fetch(1, c, s));
    if (out_len) *out_len = c.overflow ? -1 : (int64_t)c.out_len;
    if (edges_inspected) *edges_inspected = c.edges;
    if (out_queue && output) {
        size_t k = (size_t)(c.out_len < (unsigned long long)out_cap ? c.out_len : (unsigned long long)out_cap);
        GR_TRY(d_out.ToHost(out_queue, k, s));
    }
    if (edge_hits) GR_TRY(d_eh.ToHost(edge_hits, m, s));
    if (dst_hits)  GR_TRY(d_dh.ToHost(dst_hits, n, s));
    if (src_sum)   GR_TRY(d_ss.ToHost((long long *)src_sum, n, s));
    return cudaStreamSynchronize(s);
}

}  // namespace

extern "C" int grb_test_advance(grb_graph_t g, int mode, int output, const int32_t *in_queue, int64_t in_len,
                                int32_t *out_queue, int64_t out_cap, int64_t *out_len,
                                int32_t *edge_hits, int64_t *src_sum, int32_t *dst_hits, uint64_t *edges_inspected)
{
    if (!g || in_len < 0 || out_cap < 0) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    return (int)(g->g32 ? RunAdvance<int>(g->g32, g->stream, mode, output, in_queue, in_len, out_queue, out_cap, out_len,
                                          edge_hits, src_sum, dst_hits, edges_inspected)
                        : RunAdvance<long long>(g->g64, g->stream, mode, output, in_queue, in_len, out_queue, out_cap,
                                                out_len, edge_hits, src_sum, dst_hits, edges_inspected));
}

template <typename SizeT>
cudaError_t RunNeighborReduce(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t s, int op, int rtype, int atype,
                              const int32_t *in_queue, int64_t in_len, const int32_t *values, int32_t *reduced,
                              int32_t *out_queue, int64_t out_cap, int64_t *out_len)
{
    using namespace gunrock;
    typedef ProbeProblem<SizeT> Problem;
    typedef ProbeFunctor<Problem> Functor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, false> Policy;
    const size_t n = (size_t)g->nodes, m = (size_t)g->edges;
    util::DeviceArray<int> d_in, d_out, d_eh, d_dh, d_val, d_red;
    util::DeviceArray<long long> d_ss;
    oprtr::WorkProgress wp;
    oprtr::advance::DeviceInfo dev;
    GR_TRY(dev.Query(g->device));
    GR_TRY(wp.Setup(s));
    if (in_queue) GR_TRY(d_in.FromHost(in_queue, (size_t)in_len, s));
    GR_TRY(d_val.FromHost(values, rtype == 2 ? m : n, s));
    GR_TRY(d_red.Allocate((size_t)in_len + 1));
    GR_TRY(d_out.Allocate((size_t)out_cap + 1));
    GR_TRY(d_eh.Allocate(m + 1)); GR_TRY(d_eh.Zero(s));
    GR_TRY(d_dh.Allocate(n + 1)); GR_TRY(d_dh.Zero(s));
    GR_TRY(d_ss.Allocate(n + 1)); GR_TRY(d_ss.Zero(s));
    typename Problem::DataSlice ds;
    ds.edge_hits = d_eh.ptr; ds.src_sum = d_ss.ptr; ds.dst_hits = d_dh.ptr;
    ds.nodes = g->nodes; ds.local_nodes = g->local_nodes; ds.part = g->part;
    oprtr::advance::Args<int, SizeT> a;
    a.g = g->View();
    a.in_queue = in_queue ? d_in.ptr : nullptr;
    a.out_queue = out_queue ? d_out.ptr : nullptr;
    a.in_len = (unsigned long long)in_len; a.d_in_len = nullptr; a.out_cap = (unsigned long long)out_cap;
    a.slots = wp.d_slots.ptr; a.queue_index = 0; a.iteration = 7;
    a.output = out_queue ? oprtr::advance::OUT_COMPACT : oprtr::advance::OUT_NONE;
    a.type = (oprtr::advance::TYPE)(atype & 3);
    oprtr::advance::ReduceArgs<int> r;
    r.values = d_val.ptr; r.reduced = d_red.ptr;
    r.op = (oprtr::advance::REDUCE_OP)op; r.rtype = (oprtr::advance::REDUCE_TYPE)rtype;
    GR_TRY((oprtr::advance::LaunchNeighborReduce<Policy, Problem, Functor, int>(dev.sm_count, ds, a, r, s)));
    oprtr::LevelCounters c;
    GR_TRY(wp.Fetch(1, c, s));
    if (out_len) *out_len = c.overflow ? -1 : (int64_t)c.out_len;
    if (out_queue) {
        size_t k = (size_t)(c.out_len < (unsigned long long)out_cap ? c.out_len : (unsigned long long)out_cap);
        GR_TRY(d_out.ToHost(out_queue, k, s));
    }
    if (reduced) GR_TRY(d_red.ToHost(reduced, (size_t)in_len, s));
    return cudaStreamSynchronize(s);
}

extern "C" int grb_test_neighbor_reduce(grb_graph_t g, int op, int rtype, int atype, const int32_t *in_queue, int64_t in_len,
                                        const int32_t *values, int32_t *reduced, int32_t *out_queue, int64_t out_cap,
                                        int64_t *out_len)
{
    if (!g || in_len < 0 || out_cap < 0 || !values || rtype < 1 || rtype > 2) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    return (int)(g->g32 ? RunNeighborReduce<int>(g->g32, g->stream, op, rtype, atype, in_queue, in_len, values, reduced,
                                                 out_queue, out_cap, out_len)
                        : RunNeighborReduce<long long>(g->g64, g->stream, op, rtype, atype, in_queue, in_len, values,
                                                       reduced, out_queue, out_cap, out_len));
}

extern "C" int grb_test_filter(int device, int64_t nodes, int compact, const int32_t *in_queue, int64_t in_len,
                               int32_t *out_queue, int64_t out_cap, int64_t *out_len, int32_t *dst_hits)
{
    using namespace gunrock;
    if (in_len < 0 || out_cap < 0 || nodes < 0) return (int)cudaErrorInvalidValue;
    int rc = grb_select_device(device);
    if (rc) return rc;
    typedef ProbeProblem<int> Problem;
    typedef ProbeFunctor<Problem> Functor;
    typedef oprtr::filter::KernelPolicy<Problem, 256, 4, 6, 0> Policy;
    cudaStream_t s = 0;
    util::DeviceArray<int> d_in, d_out, d_dh;
    oprtr::WorkProgress wp;
    oprtr::advance::DeviceInfo dev;
    GRB_TRY(dev.Query(device));
    GRB_TRY(wp.Setup(s));
    if (in_queue) GRB_TRY(d_in.FromHost(in_queue, (size_t)in_len, s));
    GRB_TRY(d_out.Allocate((size_t)out_cap + 1));
    GRB_TRY(d_dh.Allocate((size_t)nodes + 1)); GRB_TRY(d_dh.Zero(s));
    Problem::DataSlice ds;
    ds.edge_hits = nullptr; ds.src_sum = nullptr; ds.dst_hits = d_dh.ptr;
    ds.nodes = nodes; ds.local_nodes = nodes; ds.part = graph::Partition();
    oprtr::filter::Args<int, int> fa;
    fa.in_queue = in_queue ? d_in.ptr : nullptr; fa.in_pred_queue = nullptr;
    fa.out_queue = out_queue ? d_out.ptr : nullptr;
    fa.in_len = (unsigned long long)in_len; fa.d_in_len = nullptr; fa.out_cap = (unsigned long long)out_cap;
    fa.slots = wp.d_slots.ptr; fa.queue_index = 0; fa.iteration = 3; fa.compact = compact != 0;
    GRB_TRY((oprtr::filter::LaunchKernel<Policy, Problem, Functor>(dev.sm_count, ds, fa, s)));
    oprtr::LevelCounters c;
    GRB_TRY(wp.Fetch(1, c, s));
    const unsigned long long produced = compact ? c.out_len : (unsigned long long)in_len;
    if (out_len) *out_len = c.overflow ? -1 : (int64_t)produced;
    if (out_queue) {
        size_t k = (size_t)(produced < (unsigned long long)out_cap ? produced : (unsigned long long)out_cap);
        GRB_TRY(d_out.ToHost(out_queue, k, s));
    }
    if (dst_hits) GRB_TRY(d_dh.ToHost(dst_hits, (size_t)nodes, s));
    GRB_TRY(cudaStreamSynchronize(s));
    return 0;
}

// ---- the near/far split on its own (priority_queue::Bisect; reference: gunrock/priority_queue/kernel.cuh:404-503) ----
#include "../gunrock/priority_queue/kernel.cuh"
namespace {
struct ScoreProblem {
    typedef int VertexId;
    typedef int SizeT;
    typedef int Value;
    struct DataSlice { const unsigned *score; };
};
struct ScoreFunctor {
    static __device__ __forceinline__ unsigned ComputePriorityScore(int v, const ScoreProblem::DataSlice *p) { return p->score[v]; }
};
}  // namespace

extern "C" int grb_test_bisect(int device, int64_t nodes, const uint32_t *scores, const int32_t *in_items, int64_t in_len,
                               int in_is_far, uint32_t lower, uint32_t upper, uint32_t *far_mark /* in/out, (nodes+31)/32 words */,
                               int32_t *near_out, int64_t near_cap, int64_t *near_len,
                               int32_t *far_out, int64_t far_cap, int64_t *far_len, uint32_t *min_far_score)
{
    using namespace gunrock;
    if (nodes < 0 || in_len < 0 || near_cap < 0 || far_cap < 0 || !scores || !far_mark) return (int)cudaErrorInvalidValue;
    int rc = grb_select_device(device);
    if (rc) return rc;
    cudaStream_t s = 0;
    const size_t words = (size_t)(nodes + 31) / 32;
    util::DeviceArray<unsigned> d_score, d_mark;
    util::DeviceArray<int> d_in, d_near, d_far;
    oprtr::WorkProgress wp;
    oprtr::advance::DeviceInfo dev;
    GRB_TRY(dev.Query(device));
    GRB_TRY(wp.Setup(s));
    GRB_TRY(d_score.FromHost(scores, (size_t)nodes, s));
    GRB_TRY(d_mark.FromHost(far_mark, words, s));
    GRB_TRY(d_in.FromHost(in_items, (size_t)in_len, s));
    GRB_TRY(d_near.Allocate((size_t)near_cap + 1));
    GRB_TRY(d_far.Allocate((size_t)far_cap + 1));
    ScoreProblem::DataSlice ds;
    ds.score = d_score.ptr;
    priority_queue::BisectArgs<int> a;
    a.in = d_in.ptr; a.in_len = (unsigned long long)in_len; a.d_in_len = nullptr;
    a.near_out = d_near.ptr; a.near_cap = (unsigned long long)near_cap;
    a.far_out = d_far.ptr; a.far_base = 0; a.far_cap = (unsigned long long)far_cap;
    a.far_mark = d_mark.ptr; a.in_is_far = in_is_far != 0;
    a.lower = lower; a.upper = upper;
    a.slots = wp.d_slots.ptr; a.queue_index = 0;
    GRB_TRY((priority_queue::Bisect<ScoreProblem, ScoreFunctor>(dev.sm_count, ds, a, s)));
    oprtr::LevelCounters c;
    GRB_TRY(wp.Fetch(1, c, s));
    if (near_len) *near_len = c.overflow ? -1 : (int64_t)c.out_len;
    if (far_len) *far_len = c.overflow ? -1 : (int64_t)c.aux0;
    if (min_far_score) *min_far_score = c.aux1 ? ~(uint32_t)c.aux1 : 0xffffffffu;
    if (near_out) GRB_TRY(d_near.ToHost(near_out, (size_t)(c.out_len < (unsigned long long)near_cap ? c.out_len : (unsigned long long)near_cap), s));
    if (far_out) GRB_TRY(d_far.ToHost(far_out, (size_t)(c.aux0 < (unsigned long long)far_cap ? c.aux0 : (unsigned long long)far_cap), s));
    GRB_TRY(d_mark.ToHost(far_mark, words, s));
    GRB_TRY(cudaStreamSynchronize(s));
    return 0;
}

// ---- rows binned by length (oprtr/row_bins.cuh), stand-alone ---------------------------------------------
template <typename SizeT>
static int TestRowBins(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t s, const int32_t *queue, int64_t qlen,
                       int32_t *small_rows, int32_t *medium_rows, int32_t *big_rows, int32_t *big_pieces,
                       int64_t piece_cap, int64_t *counts)
{
    using namespace gunrock;
    util::DeviceArray<int> d_queue;
    oprtr::rows::Bins<SizeT> bins;
    GRB_TRY(bins.Allocate((size_t)g->local_nodes, (size_t)g->edges));
    if (queue) GRB_TRY(d_queue.FromHost(queue, (size_t)qlen, s));
    GRB_TRY(bins.Build(g->row_offsets.ptr, queue ? d_queue.ptr : nullptr, (unsigned long long)qlen, g->part, s));
    const oprtr::rows::View &v = bins.view;
    if ((int64_t)v.n_pieces > piece_cap) return (int)cudaErrorInvalidValue;
    counts[0] = (int64_t)v.n_small; counts[1] = (int64_t)v.n_medium; counts[2] = (int64_t)v.n_big;
    counts[3] = (int64_t)v.n_pieces; counts[4] = (int64_t)v.entries;
    counts[5] = v.blocks_big; counts[6] = v.blocks_medium; counts[7] = v.blocks_small;
    GRB_TRY(bins.small_rows.ToHost(small_rows, (size_t)v.n_small, s));
    GRB_TRY(bins.medium_rows.ToHost(medium_rows, (size_t)v.n_medium, s));
    GRB_TRY(bins.big_rows.ToHost(big_rows, (size_t)v.n_big, s));
    GRB_TRY(bins.big_pieces.ToHost((int2 *)big_pieces, (size_t)v.n_pieces, s));
    GRB_TRY(cudaStreamSynchronize(s));
    return 0;
}

extern "C" int grb_test_row_bins(grb_graph_t g, const int32_t *queue, int64_t qlen, int32_t *small_rows,
                                 int32_t *medium_rows, int32_t *big_rows, int32_t *big_pieces, int64_t piece_cap,
                                 int64_t *counts)
{
    if (!g || qlen < 0 || qlen > g->local_nodes() || !small_rows || !medium_rows || !big_rows || !big_pieces || !counts)
        return (int)cudaErrorInvalidValue;
    int rc = grb_select_device(g->device());
    if (rc) return rc;
    return g->g32 ? TestRowBins<int>(g->g32, g->stream, queue, qlen, small_rows, medium_rows, big_rows, big_pieces, piece_cap, counts)
                  : TestRowBins<long long>(g->g64, g->stream, queue, qlen, small_rows, medium_rows, big_rows, big_pieces, piece_cap, counts);
}
