// bc_enactor.cuh -- the host loop of betweenness centrality (one source per Enact).
//
// Counterpart of the reference's BCEnactor::EnactBC (gunrock/app/bc/bc_enactor.cuh:208-560):
// a forward BFS that counts shortest paths (advance + filter per level) followed by a
// backward sweep that, level by level from the deepest, re-creates the level's frontier with
// a filter over all vertices and accumulates dependencies with a second advance.  Here both
// passes are the fused advance of oprtr/advance (TWC for small levels, LB otherwise); the
// forward pass appends each level to one queue, so the backward pass is one advance per
// level over a slice of that queue with no filter and no output.
#pragma once
#include <vector>
#include "bc_functor.cuh"
#include "../enactor_base.cuh"

namespace gunrock {
namespace app {
namespace bc {

template <typename Problem>
class BCEnactor : public EnactorBase {
  public:
    typedef typename Problem::VertexId  VertexId;
    typedef typename Problem::SizeT     SizeT;
    typedef typename Problem::Value     Value;
    typedef typename Problem::DataSlice DataSlice;
    typedef ForwardFunctor<VertexId, SizeT, Value, Problem>  ForwardFn;
    typedef BackwardFunctor<VertexId, SizeT, Value, Problem> BackwardFn;
    // no operator-level bitmask cull (every edge into the next level counts); out-degree of the
    // emitted vertices is tracked to size the next grid and to choose TWC / LB
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, true> AdvancePolicy;

    oprtr::advance::LbScratch<SizeT> lb_scratch;
    bool set_up = false;
    struct Level { unsigned long long start, len, edges; };
    std::vector<Level> levels;

    cudaError_t Setup(Problem *problem, cudaStream_t s)
    {
        GR_TRY(SetupBase(problem->graph->device, s));
        GR_TRY(lb_scratch.Allocate((size_t)problem->graph->local_nodes + 1, (size_t)problem->graph->edges + 1,
                                   AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        set_up = true;
        return cudaSuccess;
    }

    // reference signature: Enact<BCProblem>(context, problem, src, max_grid_size)
    cudaError_t Enact(Problem *problem, VertexId src)
    {
        if (!set_up) return util::GRError(cudaErrorNotReady, "BCEnactor::Setup not called", __FILE__, __LINE__);
        GR_TRY(ResetBase());
        levels.clear();
        typename Problem::Base::Graph &g = *problem->graph;
        if (src < 0 || (long long)src >= g.nodes) return cudaSuccess;
        const unsigned long long n = (unsigned long long)g.local_nodes;
        const unsigned osz = (unsigned)sizeof(SizeT);

        // ---- forward: BFS levels appended to one queue, sigma accumulated on the way ----------
        Level cur = {0, 1, problem->src_degree};
        VertexId depth = 0;
        while (cur.len > 0) {
            levels.push_back(cur);
            const DataSlice ds = problem->MakeDataSlice(depth);
            oprtr::advance::Args<VertexId, SizeT> a = LevelArgs(problem, cur, depth);
            a.out_queue = problem->level_queue.ptr + cur.start + cur.len;
            a.out_cap = n - (cur.start + cur.len);
            a.output = oprtr::advance::OUT_COMPACT;
            GR_TRY(Launch<ForwardFn>(ds, a, cur.edges));
            oprtr::LevelCounters c;
            GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, c, stream));
            if (c.overflow)
                return util::GRError(cudaErrorInvalidConfiguration, "BC level queue overflow", __FILE__, __LINE__);
            enactor_stats.total_queued += cur.len;
            enactor_stats.edges_inspected += c.edges;
            enactor_stats.algo_bytes += cur.len * (4 + 2 * osz) + 12 * c.edges + 4 * c.out_len;
            frontier_attribute.queue_index++;
            ++enactor_stats.iteration;
            Level next = {cur.start + cur.len, c.out_len, c.out_degree};
            cur = next;
            ++depth;
        }
        // ---- backward: dependencies, deepest level first (the deepest has no successors; the
        // source, level 0, accumulates nothing: bc_functor.cuh:195) ----------------------------------
        for (long long l = (long long)levels.size() - 2; l >= 1; --l) {
            const DataSlice ds = problem->MakeDataSlice((VertexId)l);
            oprtr::advance::Args<VertexId, SizeT> a = LevelArgs(problem, levels[l], (VertexId)l);
            a.out_queue = nullptr; a.out_cap = 0;
            a.output = oprtr::advance::OUT_NONE;
            GR_TRY(Launch<BackwardFn>(ds, a, levels[l].edges));
            oprtr::LevelCounters c;
            GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, c, stream));
            enactor_stats.edges_inspected += c.edges;
            enactor_stats.algo_bytes += levels[l].len * (4 + 2 * osz) + 20 * c.edges;
            frontier_attribute.queue_index++;
            ++enactor_stats.iteration;
        }
        return cudaSuccess;
    }

  private:
    oprtr::advance::Args<VertexId, SizeT> LevelArgs(Problem *problem, const Level &lv, VertexId depth)
    {
        oprtr::advance::Args<VertexId, SizeT> a;
        a.g = problem->graph->View();
        a.in_queue = problem->level_queue.ptr + lv.start;
        a.in_len = lv.len; a.d_in_len = nullptr;
        a.slots = work_progress.d_slots.ptr;
        a.queue_index = frontier_attribute.queue_index;
        a.iteration = depth;
        return a;
    }

    template <typename Functor>
    cudaError_t Launch(const DataSlice &ds, const oprtr::advance::Args<VertexId, SizeT> &a, unsigned long long edges)
    {
        const oprtr::advance::MODE mode = edges >= 4096 ? oprtr::advance::LB : oprtr::advance::TWC_FORWARD;
        enactor_stats.launches += (mode == oprtr::advance::LB) ? 2 : 1;
        return oprtr::advance::LaunchKernel<AdvancePolicy, Problem, Functor>(mode, dev, ds, a, &lb_scratch, edges, stream);
    }
};

}  // namespace bc
}  // namespace app
}  // namespace gunrock
