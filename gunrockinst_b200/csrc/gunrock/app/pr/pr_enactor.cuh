// pr_enactor.cuh -- the host loop of PageRank.
//
// Counterpart of the reference's PREnactor::EnactPR (gunrock/app/pr/pr_enactor.cuh:186-512):
//   (a) prune zero-out-degree vertices: advance(RemoveZeroDegreeNodeFunctor) + filter + copy
//       of the reduced degrees, repeated until the queue length is stable (:247-329);
//   (b) power iteration over the surviving queue: advance(PRFunctor) with no output queue
//       (:346-370), filter(PRFunctor) with no output queue whose survivor count is the number
//       of vertices still moving (:405-430), curr <- next, next <- 0 (:444-451), until that
//       count is 0 or max_iter sweeps ran (:472-474);
//   (c) sort (rank, id) descending (:497-500) -- done by Problem::Extract here.
// The reference pays per sweep 4-5 advance launches, a filter, two grid-128 copy kernels and
// three host round trips; here a sweep is ScanKernel + ExpandKernel + filter + one memset, the
// curr/next exchange is a pointer swap, and one 64-byte counter fetch decides termination.
//
// Algorithmic HBM bytes per sweep: queue 4 + offsets
// 2*sizeof(SizeT) per vertex, per edge 4 (column) + 4 (degree probe) + 8 (contribution) + 8
// (atomic add), per vertex 32 (next, curr, contrib, degree) in the filter.
#pragma once
#include "pr_functor.cuh"
#include "../enactor_base.cuh"

namespace gunrock {
namespace app {
namespace pr {

template <typename Problem>
class PREnactor : public EnactorBase {
  public:
    typedef typename Problem::VertexId  VertexId;
    typedef typename Problem::SizeT     SizeT;
    typedef typename Problem::Value     Value;
    typedef typename Problem::DataSlice DataSlice;
    typedef PRFunctor<VertexId, SizeT, Value, Problem>                   PrFunctor;
    typedef RemoveZeroDegreeNodeFunctor<VertexId, SizeT, Value, Problem> RemoveZeroFunctor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, false> AdvancePolicy;
    typedef oprtr::filter::KernelPolicy<Problem, 256, 4, 6, 0>             FilterPolicy;

    oprtr::advance::LbScratch<SizeT> lb_scratch;
    bool set_up = false;
    int  prune_sweeps = 0;
    unsigned long long valid_nodes = 0;
    float sweep_ms = 0;                      // CUDA-event time of the power iteration alone (:343-345,478-480)
    cudaEvent_t ev_a = nullptr, ev_b = nullptr;

    ~PREnactor()
    {
        if (ev_a) cudaEventDestroy(ev_a);
        if (ev_b) cudaEventDestroy(ev_b);
    }

    cudaError_t Setup(Problem *problem, cudaStream_t s)
    {
        GR_TRY(SetupBase(problem->graph->device, s));
        GR_TRY(lb_scratch.Allocate((size_t)problem->graph->local_nodes + 1, (size_t)problem->graph->edges + 1,
                                   AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        if (!ev_a) { GR_TRY(cudaEventCreate(&ev_a)); GR_TRY(cudaEventCreate(&ev_b)); }
        set_up = true;
        return cudaSuccess;
    }

    // reference signature: Enact<PRProblem>(context, problem, max_iteration, max_grid_size)
    cudaError_t Enact(Problem *problem, long long max_iteration)
    {
        if (!set_up) return util::GRError(cudaErrorNotReady, "PREnactor::Setup not called", __FILE__, __LINE__);
        GR_TRY(ResetBase());
        prune_sweeps = 0; valid_nodes = 0; sweep_ms = 0;
        typename Problem::Base::Graph &g = *problem->graph;
        const unsigned long long n = (unsigned long long)g.local_nodes;
        if (n == 0) return cudaSuccess;
        const unsigned osz = (unsigned)sizeof(SizeT);
        const unsigned long long m = (unsigned long long)g.edges;

        // ---- (a) prune vertices without out-edges -----------------------------------------------------
        const VertexId *in_queue = nullptr;                  // identity frontier first
        unsigned long long qlen = n, prev = 0;
        while (prev != qlen) {
            prev = qlen;
            DataSlice ds = problem->MakeDataSlice();
            oprtr::advance::Args<VertexId, SizeT> a = AdvArgs(problem, in_queue, qlen);
            GR_TRY((oprtr::advance::LaunchKernel<AdvancePolicy, Problem, RemoveZeroFunctor>(
                oprtr::advance::LB, dev, ds, a, &lb_scratch, m, stream)));
            frontier_attribute.queue_index++;
            oprtr::filter::Args<VertexId, Value> fa = FilterArgs(in_queue, qlen);
            fa.out_queue = problem->frontier_queues[frontier_attribute.selector ^ 1].ptr;
            fa.out_cap = problem->frontier_elements[frontier_attribute.selector ^ 1];
            GR_TRY((oprtr::filter::LaunchKernel<FilterPolicy, Problem, RemoveZeroFunctor>(dev.sm_count, ds, fa, stream)));
            GR_TRY(cudaMemcpyAsync(problem->degrees.ptr, problem->degrees_pong.ptr, sizeof(int) * (size_t)n,
                                   cudaMemcpyDeviceToDevice, stream));
            enactor_stats.launches += 3;
            oprtr::LevelCounters c;
            GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, c, stream));
            if (c.overflow) return util::GRError(cudaErrorInvalidConfiguration, "PR queue overflow", __FILE__, __LINE__);
            frontier_attribute.queue_index++;
            frontier_attribute.selector ^= 1;
            in_queue = problem->frontier_queues[frontier_attribute.selector].ptr;
            qlen = c.out_len;
            ++prune_sweeps;
        }
        valid_nodes = qlen;
        if (qlen == 0 || max_iteration <= 0) return cudaSuccess;
        InitContribKernel<<<util::FillGrid(n), 256, 0, stream>>>(problem->degrees.ptr, problem->rank[problem->curr].ptr,
                                                                problem->contrib.ptr, n);
        GR_TRY(cudaGetLastError());

        // ---- (b) power iteration ----------------------------------------------------------------------------
        GR_TRY(cudaEventRecord(ev_a, stream));
        for (;;) {
            DataSlice ds = problem->MakeDataSlice();
            oprtr::advance::Args<VertexId, SizeT> a = AdvArgs(problem, in_queue, qlen);
            GR_TRY((oprtr::advance::LaunchKernel<AdvancePolicy, Problem, PrFunctor>(
                oprtr::advance::LB, dev, ds, a, &lb_scratch, m, stream)));
            frontier_attribute.queue_index++;
            oprtr::filter::Args<VertexId, Value> fa = FilterArgs(in_queue, qlen);      // survivors are only counted
            GR_TRY((oprtr::filter::LaunchKernel<FilterPolicy, Problem, PrFunctor>(dev.sm_count, ds, fa, stream)));
            // curr <- next is a pointer swap; the new next (old curr) starts from zero
            problem->curr ^= 1;
            GR_TRY(cudaMemsetAsync(problem->rank[problem->curr ^ 1].ptr, 0, sizeof(Value) * (size_t)n, stream));
            enactor_stats.launches += 3;
            oprtr::LevelCounters ca, cf;
            GR_TRY(work_progress.Fetch(frontier_attribute.queue_index, ca, stream));
            GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, cf, stream));
            frontier_attribute.queue_index++;
            ++enactor_stats.iteration;
            enactor_stats.total_queued += qlen;
            enactor_stats.edges_inspected += ca.edges;
            enactor_stats.algo_bytes += qlen * (4ull + 2ull * osz) + 24ull * ca.edges + 32ull * qlen + 8ull * n;
            if (cf.out_len == 0 || enactor_stats.iteration >= max_iteration) break;
        }
        GR_TRY(cudaEventRecord(ev_b, stream));
        GR_TRY(cudaEventSynchronize(ev_b));
        GR_TRY(cudaEventElapsedTime(&sweep_ms, ev_a, ev_b));
        return cudaSuccess;
    }

  private:
    oprtr::advance::Args<VertexId, SizeT> AdvArgs(Problem *problem, const VertexId *in_queue, unsigned long long qlen)
    {
        oprtr::advance::Args<VertexId, SizeT> a;
        a.g = problem->graph->View();
        a.in_queue = in_queue; a.out_queue = nullptr;       // functor side effects only (:363)
        a.in_len = qlen; a.d_in_len = nullptr; a.out_cap = 0;
        a.slots = work_progress.d_slots.ptr;
        a.queue_index = frontier_attribute.queue_index;
        a.iteration = (long long)enactor_stats.iteration;
        a.output = oprtr::advance::OUT_NONE;
        return a;
    }
    oprtr::filter::Args<VertexId, Value> FilterArgs(const VertexId *in_queue, unsigned long long qlen)
    {
        oprtr::filter::Args<VertexId, Value> fa;
        fa.in_queue = in_queue; fa.in_pred_queue = nullptr; fa.out_queue = nullptr;
        fa.in_len = qlen; fa.d_in_len = nullptr; fa.out_cap = 0;
        fa.slots = work_progress.d_slots.ptr;
        fa.queue_index = frontier_attribute.queue_index;
        fa.iteration = (long long)enactor_stats.iteration;
        fa.compact = true;
        return fa;
    }
};

}  // namespace pr
}  // namespace app
}  // namespace gunrock
