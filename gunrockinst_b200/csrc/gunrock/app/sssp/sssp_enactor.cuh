// sssp_enactor.cuh -- the host loop of single-source shortest paths.
//
// Counterpart of the reference's SSSPEnactor::EnactSSSP (gunrock/app/sssp/sssp_enactor.cuh:
// 203-464): advance (relax) -> filter -> priority_queue::Bisect into a near queue (next
// frontier) and a far pile, bumping the bucket level when the near queue runs dry (:384-420).
// The reference spends per iteration 4-5 advance launches, a filter, five queue kernels,
// four library scans and five blocking 4-byte copies; here an iteration is
//     memset(next-frontier bits)  +  ScanKernel + ExpandKernel  +  BisectKernel
// and two 64-byte counter fetches.  Duplicates never reach a queue (test-and-set bits in
// the functor / the pile), so both queues and both piles hold at most n ids.
//
// Result: exact shortest distances (fixed point of atomicMin relaxation, independent of the
// bucket width), parents that always realise the distance next to them.
#pragma once
#include "sssp_functor.cuh"
#include "../enactor_base.cuh"

namespace gunrock {
namespace app {
namespace sssp {

template <typename Problem>
class SSSPEnactor : public EnactorBase {
  public:
    typedef typename Problem::VertexId  VertexId;
    typedef typename Problem::SizeT     SizeT;
    typedef typename Problem::DataSlice DataSlice;
    typedef SSSPFunctor<VertexId, SizeT, Problem> SsspFunctor;
    typedef PQFunctor<VertexId, SizeT, Problem>   PqFunctor;
    // no bitmask pre-cull (a visited vertex can still improve); out-degree of the emitted
    // vertices is tracked to size the next grid
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, true> AdvancePolicy;

    oprtr::advance::LbScratch<SizeT> lb_scratch;
    bool set_up = false;
    unsigned long long relaxations = 0;       // neighbour entries inspected
    unsigned long long bisected = 0;          // items that went through the priority queue
    int bucket_jumps = 0;

    cudaError_t Setup(Problem *problem, cudaStream_t s)
    {
        GR_TRY(SetupBase(problem->graph->device, s));
        GR_TRY(lb_scratch.Allocate((size_t)problem->graph->local_nodes + 1, (size_t)problem->graph->edges + 1,
                                   AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        set_up = true;
        return cudaSuccess;
    }

    // reference signature: Enact<SSSPProblem>(context, problem, src, queue_sizing, max_grid_size, traversal_mode)
    cudaError_t Enact(Problem *problem, VertexId src, int traversal_mode = 0)
    {
        if (!set_up) return util::GRError(cudaErrorNotReady, "SSSPEnactor::Setup not called", __FILE__, __LINE__);
        GR_TRY(ResetBase());
        relaxations = bisected = 0; bucket_jumps = 0;
        typename Problem::Base::Graph &g = *problem->graph;
        if (src < 0 || (long long)src >= g.nodes) return cudaSuccess;

        const DataSlice ds = problem->MakeDataSlice();
        auto &pile = problem->pile;
        const unsigned long long n = (unsigned long long)g.local_nodes;
        const size_t mark_bytes = sizeof(unsigned) * problem->mark.size;
        unsigned long long near_len = 1, near_edges = problem->src_degree;
        unsigned level = 0;                     // current bucket: distances in [level*delta, (level+1)*delta)
        const unsigned osz = (unsigned)sizeof(SizeT);

        for (;;) {
            while (near_len > 0) {
                // ---- relax every out-edge of the near queue --------------------------------------
                GR_TRY(cudaMemsetAsync(problem->mark.ptr, 0, mark_bytes, stream));
                oprtr::advance::Args<VertexId, SizeT> a;
                a.g = g.View();
                a.in_queue  = problem->frontier_queues[frontier_attribute.selector].ptr;
                a.out_queue = problem->frontier_queues[frontier_attribute.selector ^ 1].ptr;
                a.in_len = near_len; a.d_in_len = nullptr;
                a.out_cap = problem->frontier_elements[frontier_attribute.selector ^ 1];
                a.slots = work_progress.d_slots.ptr;
                a.queue_index = frontier_attribute.queue_index;
                a.iteration = (long long)enactor_stats.iteration;
                a.output = oprtr::advance::OUT_COMPACT;
                oprtr::advance::MODE mode = (traversal_mode == 2) ? oprtr::advance::TWC_FORWARD
                                          : (traversal_mode == 1 || near_edges >= 4096) ? oprtr::advance::LB
                                          : oprtr::advance::TWC_FORWARD;
                GR_TRY((oprtr::advance::LaunchKernel<AdvancePolicy, Problem, SsspFunctor>(
                    mode, dev, ds, a, &lb_scratch, near_edges, stream)));
                enactor_stats.launches += (mode == oprtr::advance::LB) ? 2 : 1;
                oprtr::LevelCounters c;
                GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, c, stream));
                if (c.overflow)
                    return util::GRError(cudaErrorInvalidConfiguration, "frontier queue overflow", __FILE__, __LINE__);
                enactor_stats.total_queued += near_len;
                relaxations += c.edges;
                enactor_stats.algo_bytes += near_len * (4 + 2 * osz + 4) + 12 * c.edges + 4 * c.out_len;
                frontier_attribute.queue_index++;
                frontier_attribute.selector ^= 1;
                ++enactor_stats.iteration;
                const unsigned long long out_len = c.out_len;
                if (out_len == 0) { near_len = 0; break; }

                // ---- split what improved into near (this bucket) and far ---------------------------
                priority_queue::BisectArgs<VertexId> b;
                b.in = problem->frontier_queues[frontier_attribute.selector].ptr;
                b.in_len = out_len; b.d_in_len = nullptr;
                b.near_out = problem->frontier_queues[frontier_attribute.selector ^ 1].ptr;
                b.near_cap = problem->frontier_elements[frontier_attribute.selector ^ 1];
                b.far_out = pile.far[pile.selector].ptr;
                b.far_base = pile.far_len; b.far_cap = pile.far[pile.selector].cap;
                b.far_mark = pile.far_mark.ptr; b.in_is_far = false;
                b.lower = 0; b.upper = level + 1;
                b.slots = work_progress.d_slots.ptr;
                b.queue_index = frontier_attribute.queue_index;
                GR_TRY((priority_queue::Bisect<Problem, PqFunctor>(dev.sm_count, ds, b, stream)));
                ++enactor_stats.launches;
                oprtr::LevelCounters cb;
                GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, cb, stream));
                if (cb.overflow)
                    return util::GRError(cudaErrorInvalidConfiguration, "near/far pile overflow", __FILE__, __LINE__);
                bisected += out_len;
                enactor_stats.algo_bytes += 8 * out_len + 4 * (cb.out_len + cb.aux0);
                frontier_attribute.queue_index++;
                frontier_attribute.selector ^= 1;
                near_len = cb.out_len;
                pile.far_len += cb.aux0;
                near_edges = c.out_degree;      // upper bound: the near queue is a subset of what was emitted
            }
            if (pile.far_len == 0) break;

            // ---- near queue dry: open the next non-empty bucket of the far pile ----------------------
            // first pass finds the smallest bucket present, second pass moves it to the near queue
            unsigned next_level = level + 1;
            for (int pass = 0; pass < 2; ++pass) {
                priority_queue::BisectArgs<VertexId> b;
                b.in = pile.far[pile.selector].ptr;
                b.in_len = pile.far_len; b.d_in_len = nullptr;
                b.near_out = problem->frontier_queues[frontier_attribute.selector].ptr;
                b.near_cap = problem->frontier_elements[frontier_attribute.selector];
                b.far_out = pile.far[pile.selector ^ 1].ptr;
                b.far_base = 0; b.far_cap = pile.far[pile.selector ^ 1].cap;
                b.far_mark = pile.far_mark.ptr; b.in_is_far = true;
                // entries below the bucket just finished were served from a near queue already
                b.lower = level + 1; b.upper = next_level + 1;
                b.slots = work_progress.d_slots.ptr;
                b.queue_index = frontier_attribute.queue_index;
                GR_TRY((priority_queue::Bisect<Problem, PqFunctor>(dev.sm_count, ds, b, stream)));
                ++enactor_stats.launches;
                oprtr::LevelCounters cb;
                GR_TRY(work_progress.Fetch(frontier_attribute.queue_index + 1, cb, stream));
                if (cb.overflow)
                    return util::GRError(cudaErrorInvalidConfiguration, "near/far pile overflow", __FILE__, __LINE__);
                bisected += pile.far_len;
                enactor_stats.algo_bytes += 8 * pile.far_len + 4 * (cb.out_len + cb.aux0);
                frontier_attribute.queue_index++;
                pile.selector ^= 1;
                pile.far_len = cb.aux0;
                near_len = cb.out_len;
                if (near_len > 0 || pile.far_len == 0) break;
                // nothing in bucket level+1: jump straight to the smallest bucket that is present
                next_level = ~(unsigned)cb.aux1;
                ++bucket_jumps;
            }
            level = next_level;
            near_edges = (unsigned long long)g.edges;   // unknown: size the grid for the worst case
            if (near_len == 0 && pile.far_len == 0) break;
        }
        return cudaSuccess;
    }
};

}  // namespace sssp
}  // namespace app
}  // namespace gunrock
