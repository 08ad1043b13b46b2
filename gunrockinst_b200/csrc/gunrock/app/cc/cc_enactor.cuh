// cc_enactor.cuh -- the host side of connected components.
//
// Counterpart of the reference's CCEnactor::EnactCC (gunrock/app/cc/cc_enactor.cuh:160-862):
// there, a do/while of filter launches (HookInit, then rounds of PtrJump until a device flag
// settles, UpdateMask, HookMax, PtrJumpMask/Unmask), each round ending in a blocking flag
// copy -- O(10^2) launches and as many host round trips on an R-MAT graph.
//
// Here the whole computation is TWO operator launches and no host decision in between:
//   1. advance (LB mode, identity frontier = every vertex, OUT_NONE) with HookFunctor:
//      every CSR entry chases and links the roots of its end points;
//   2. filter in apply-only mode (the reference's Kernel2) with PtrJumpFunctor over the
//      identity frontier: flatten every vertex to its root and count the roots.
// Labels: smallest vertex id of the component, as in the reference.
// Default route of pass 1 (cc_row_hook.cuh): the rows binned by length inside Enact (one small
// kernel + one 64-byte read-back, part of the measured time) and ONE kernel in which the lanes
// that walk a row keep its root (resident grid with tickets; GRB_CC_HOOK=rows1: static grid);
// GRB_CC_HOOK=lb selects the advance.
//
// Algorithmic HBM bytes: hook pass 2*sizeof(SizeT) per vertex + 4 (column id) + 8 (two
// parent probes) per CSR entry; jump pass 12 per vertex.
#pragma once
#include <cstdlib>
#include <cstring>
#include "cc_functor.cuh"
#include "cc_row_hook.cuh"
#include "../enactor_base.cuh"

namespace gunrock {
namespace app {
namespace cc {

template <typename Problem>
class CCEnactor : public EnactorBase {
  public:
    typedef typename Problem::VertexId  VertexId;
    typedef typename Problem::SizeT     SizeT;
    typedef typename Problem::Value     Value;
    typedef typename Problem::DataSlice DataSlice;
    typedef HookFunctor<VertexId, SizeT, Value, Problem>    HookFn;
    typedef PtrJumpFunctor<VertexId, SizeT, Value, Problem> JumpFn;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, false, false> AdvancePolicy;
    typedef oprtr::filter::KernelPolicy<Problem, 256, 4, 6, 0>             FilterPolicy;

    oprtr::advance::LbScratch<SizeT> lb_scratch;
    oprtr::rows::Bins<SizeT> row_bins;
    bool by_rows = false, row_tickets = false;
    util::DeviceArray<unsigned long long> hook_ticket;
    int ticket_ctas_per_sm = 1;
    bool set_up = false;

    cudaError_t Setup(Problem *problem, cudaStream_t s)
    {
        GR_TRY(SetupBase(problem->graph->device, s));
        const char *route = getenv("GRB_CC_HOOK");
        by_rows = problem->graph->part.log_parts == 0 && !(route && strcmp(route, "lb") == 0);
        row_tickets = by_rows && !(route && strcmp(route, "rows1") == 0);      // rows1: the static grid of CcRowHookKernel
        if (by_rows) {
            GR_TRY(row_bins.Allocate((size_t)problem->graph->local_nodes, (size_t)problem->graph->edges));
            GR_TRY(hook_ticket.Allocate(1));
            GR_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ticket_ctas_per_sm, CcRowHookTicketKernel<SizeT>,
                                                                 oprtr::rows::kThreads, 0));
            if (ticket_ctas_per_sm < 1) ticket_ctas_per_sm = 1;
        }
        GR_TRY(lb_scratch.Allocate((size_t)problem->graph->local_nodes + 1, (size_t)problem->graph->edges + 1,
                                   AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        set_up = true;
        return cudaSuccess;
    }

    // reference signature: Enact<CCProblem>(problem, max_grid_size)
    cudaError_t Enact(Problem *problem)
    {
        if (!set_up) return util::GRError(cudaErrorNotReady, "CCEnactor::Setup not called", __FILE__, __LINE__);
        GR_TRY(ResetBase());
        typename Problem::Base::Graph &g = *problem->graph;
        const unsigned long long n = (unsigned long long)g.local_nodes;
        if (n == 0) return cudaSuccess;
        const DataSlice ds = problem->MakeDataSlice();
        const unsigned osz = (unsigned)sizeof(SizeT);

        // ---- 1. hook: one pass over the CSR entries ------------------------------------------------
        unsigned long long row_entries = 0;
        if (by_rows) {
            GR_TRY(row_bins.Build(g.row_offsets.ptr, nullptr, n, g.part, stream));
            CcRowHookArgs<SizeT> h;
            h.row_offsets = g.row_offsets.ptr; h.col_indices = g.col_indices.ptr;
            h.parent = ds.d_component_ids; h.skip_upper = ds.skip_upper;
            h.bins = row_bins.view;
            row_entries = h.bins.entries;
            if (h.bins.Blocks() && row_tickets) {
                const unsigned long long units = h.bins.n_pieces + (h.bins.n_medium + kHookMediumUnit - 1) / kHookMediumUnit +
                                                 (h.bins.n_small + kHookSmallUnit - 1) / kHookSmallUnit;
                const unsigned long long want = (units + oprtr::rows::kWarps - 1) / oprtr::rows::kWarps;
                const unsigned long long resident = (unsigned long long)dev.sm_count * (unsigned long long)ticket_ctas_per_sm;
                GR_TRY(cudaMemsetAsync(hook_ticket.ptr, 0, sizeof(unsigned long long), stream));
                CcRowHookTicketKernel<SizeT><<<(unsigned)(want < resident ? want : resident), oprtr::rows::kThreads, 0, stream>>>(
                    h, hook_ticket.ptr);
                GR_TRY(cudaGetLastError());
            } else if (h.bins.Blocks()) {
                CcRowHookKernel<SizeT><<<h.bins.Blocks(), oprtr::rows::kThreads, 0, stream>>>(h);
                GR_TRY(cudaGetLastError());
            }
            enactor_stats.launches += 2;
            frontier_attribute.queue_index++;
        } else {
            oprtr::advance::Args<VertexId, SizeT> a;
            a.g = g.View();
            a.in_queue = nullptr; a.out_queue = nullptr;        // identity frontier, functor side effects only
            a.in_len = n; a.d_in_len = nullptr; a.out_cap = 0;
            a.slots = work_progress.d_slots.ptr;
            a.queue_index = frontier_attribute.queue_index;
            a.iteration = 0;
            a.output = oprtr::advance::OUT_NONE;
            GR_TRY((oprtr::advance::LaunchKernel<AdvancePolicy, Problem, HookFn>(
                oprtr::advance::LB, dev, ds, a, &lb_scratch, (unsigned long long)g.edges, stream)));
            enactor_stats.launches += 2;
            frontier_attribute.queue_index++;
        }

        // ---- 2. pointer jumping: flatten + count roots ---------------------------------------------
        if (by_rows) {
            CcFlattenKernel<<<util::FillGrid(n, 256, (unsigned)dev.sm_count * 8u), 256, 0, stream>>>(
                ds.d_component_ids, (long long)n, ds.d_num_components);
            GR_TRY(cudaGetLastError());
        } else {
            oprtr::filter::Args<VertexId, Value> fa;
            fa.in_queue = nullptr; fa.in_pred_queue = nullptr; fa.out_queue = nullptr;
            fa.in_len = n; fa.d_in_len = nullptr; fa.out_cap = 0;
            fa.slots = work_progress.d_slots.ptr;
            fa.queue_index = frontier_attribute.queue_index;
            fa.iteration = 1;
            fa.compact = false;
            GR_TRY((oprtr::filter::LaunchKernel<FilterPolicy, Problem, JumpFn>(dev.sm_count, ds, fa, stream)));
        }
        ++enactor_stats.launches;
        frontier_attribute.queue_index++;

        GR_TRY(cudaMemcpyAsync(&problem->num_components, problem->counter.ptr, sizeof(unsigned long long),
                               cudaMemcpyDeviceToHost, stream));
        oprtr::LevelCounters c;
        GR_TRY(work_progress.Fetch(1, c, stream));           // the hook pass accumulated into slot 1
        if (by_rows) c.edges = row_entries;
        enactor_stats.edges_inspected = c.edges;
        enactor_stats.total_queued = 2 * n;
        enactor_stats.algo_bytes = n * (2ull * osz) + 12ull * c.edges + 12ull * n;
        enactor_stats.iteration = 2;
        return cudaSuccess;
    }
};

}  // namespace cc
}  // namespace app
}  // namespace gunrock
