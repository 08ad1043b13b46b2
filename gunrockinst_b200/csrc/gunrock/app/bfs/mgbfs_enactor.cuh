// mgbfs_enactor.cuh -- per-level pieces of a 1D vertex-partitioned BFS, one device per process.
//
// The reference declares multi-GPU state (gunrock/app/problem_base.cuh:186-210: owner = v % P,
// local row = v / P) but never implements it (:336-338 "TODO").  This enactor supplies the
// device side of that missing path; the host loop and the collectives live in the harness
// (gunrockinst_b200/partitioned.py: torch.distributed / NCCL over NVLink), which calls the
// steps below on the graph's stream between its collectives.
//
// Every device holds the rows it owns (global neighbour ids) and three GLOBAL 1-bit-per-vertex
// maps laid out owner-major, so that a device's bits are one contiguous, word-aligned segment:
//     visited        own segment authoritative, other segments are cull hints
//     frontier map   x2, own segment written by this device, all-gathered before a pull level
//     next map       discoveries of a push level, whoever owns them
// A level is
//   push:  Push()   advance over the local queue; every unvisited neighbour sets its bit in
//                   the next map (no queue output, no labels yet)
//          --- all-to-all of the next map's segments (harness) ---
//          Absorb() OR of the received segments & ~visited -> labels, own frontier segment,
//                   local queue, counters
//   pull:  --- all-gather of the frontier map (harness) ---
//          Pull()   bottom-up sweep of the owned rows against the global frontier map
// Counters of a level (oprtr::LevelCounters, 64 bytes) sit at a fixed device address that the
// harness all-gathers, so one collective + one host read per level carries the frontier
// length, its degree sum and the edges inspected of every device.
#pragma once
#include "bfs_enactor.cuh"

namespace gunrock {
namespace app {
namespace bfs {

// push-level functor: mark the neighbour in the dense discovery map
template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct MgPushFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        const unsigned long long pos = problem->BitPos(d_id);
        const unsigned bit = 1u << (pos & 31);
        unsigned *word = problem->d_next_map + (pos >> 5);
        if (!(util::LoadCA(word) & bit)) atomicOr(word, bit);
        return false;                                   // nothing is enqueued before the exchange
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0) {}
    static __device__ __forceinline__ bool CondFilter(VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) {}
};

// OR the `parts` received copies of this device's segment, keep what is new, label it, publish
// it as this device's segment of the frontier map and enqueue it (CTA-aggregated)
template <typename VertexId, typename SizeT, typename DataSlice, int THREADS>
__global__ void __launch_bounds__(THREADS)
AbsorbDenseKernel(const unsigned *recv, int parts, unsigned *visited, unsigned *frontier_out, VertexId *queue,
                  unsigned long long cap, const SizeT *row_offsets, oprtr::LevelCounters *slot, DataSlice ds)
{
    __shared__ unsigned s_warp[THREADS / 32 + 1];
    __shared__ unsigned long long s_red[THREADS / 32 + 1];
    __shared__ unsigned long long s_base;
    const long long words = (ds.local_nodes + 31) >> 5;
    const long long seg = ((long long)ds.part.rank * ds.local_nodes) >> 5;
    const long long rounds = (words + (long long)gridDim.x * THREADS - 1) / ((long long)gridDim.x * THREADS);
    unsigned long long my_degree = 0;
    for (long long r = 0; r < rounds; ++r) {
        const long long w = (r * gridDim.x + blockIdx.x) * THREADS + threadIdx.x;
        unsigned bits = 0;
        if (w < words) {
            unsigned acc = 0;
            for (int p = 0; p < parts; ++p) acc |= util::LoadStream(recv + (long long)p * words + w);
            const unsigned vis = visited[seg + w];
            bits = acc & ~vis;
            frontier_out[seg + w] = bits;
            if (bits) visited[seg + w] = vis | bits;
        }
        unsigned total;
        unsigned excl = util::BlockExclusiveSum<THREADS>((unsigned)__popc(bits), s_warp, total);
        if (total == 0) continue;                       // uniform
        if (threadIdx.x == 0) s_base = atomicAdd(&slot->out_len, (unsigned long long)total);
        __syncthreads();
        unsigned long long at = s_base + excl;
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
    oprtr::advance::BlockAccumulate<THREADS>(my_degree, s_red, &slot->out_degree);
}

// visited |= map over the whole global bitmap (refreshes the cull hints after an all-gather)
static __global__ void MergeVisitedKernel(unsigned *visited, const unsigned *map, unsigned long long words)
{
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long w = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += stride) {
        const unsigned m = map[w];
        if (m) visited[w] |= m;
    }
}

template <typename Problem>
class MGBFSEnactor : public EnactorBase {
  public:
    typedef typename Problem::VertexId VertexId;
    typedef typename Problem::SizeT    SizeT;
    typedef typename Problem::Value    Value;
    typedef typename Problem::DataSlice DataSlice;
    typedef MgPushFunctor<VertexId, SizeT, Value, Problem>     PushFunctor;
    typedef ReverseBFSFunctor<VertexId, SizeT, Value, Problem> PullFunctor;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 5, true, false>  AdvancePolicy;
    typedef oprtr::advance::KernelPolicy<Problem, 256, 4, 4, false, false> PullPolicy;

    oprtr::advance::LbScratch<SizeT> lb_scratch;
    int in_sel = 0;                      // frontier_queues[in_sel] is the local frontier
    bool set_up = false;

    oprtr::LevelCounters *Counters() { return work_progress.Slot(1); }

    cudaError_t Setup(Problem *problem, cudaStream_t s)
    {
        GR_TRY(SetupBase(problem->graph->device, s));
        GR_TRY(lb_scratch.Allocate((size_t)problem->graph->local_nodes + 1, (size_t)problem->graph->edges + 1,
                                   AdvancePolicy::SCAN_TILE, AdvancePolicy::TILE_EDGES, s));
        GR_TRY(problem->PreparePull());
        GR_TRY(problem->PrepareExchange());
        set_up = true;
        return cudaSuccess;
    }
    cudaError_t Reset(Problem *problem)
    {
        in_sel = 0;
        GR_TRY(ResetBase());
        GR_TRY(problem->frontier_map[0].Zero(stream));
        GR_TRY(problem->frontier_map[1].Zero(stream));
        GR_TRY(problem->next_map.Zero(stream));
        return problem->recv_map.Zero(stream);
    }
    cudaError_t ZeroCounters() { return cudaMemsetAsync(Counters(), 0, sizeof(oprtr::LevelCounters), stream); }

    // push level, part 1: expand the local queue into the dense discovery map
    cudaError_t Push(Problem *problem, VertexId level, unsigned long long queue_len, unsigned long long edges_hint)
    {
        GR_TRY(ZeroCounters());
        GR_TRY(problem->next_map.Zero(stream));
        if (queue_len == 0) return cudaSuccess;
        DataSlice ds = problem->MakeDataSlice(level);
        oprtr::advance::Args<VertexId, SizeT> a;
        a.g = problem->graph->View();
        a.in_queue = problem->frontier_queues[in_sel].ptr; a.out_queue = nullptr;
        a.in_len = queue_len; a.d_in_len = nullptr; a.out_cap = 0;
        a.slots = work_progress.d_slots.ptr; a.queue_index = 0; a.iteration = level;
        a.output = oprtr::advance::OUT_NONE;
        oprtr::advance::MODE mode = edges_hint >= 4096 ? oprtr::advance::LB : oprtr::advance::TWC_FORWARD;
        enactor_stats.launches += (mode == oprtr::advance::LB) ? 2 : 1;
        return oprtr::advance::LaunchKernel<AdvancePolicy, Problem, PushFunctor>(mode, dev, ds, a, &lb_scratch,
                                                                                 edges_hint, stream);
    }
    // push level, part 2 (after the all-to-all of the next map into recv_map)
    cudaError_t Absorb(Problem *problem, VertexId level, int cur_out)
    {
        DataSlice ds = problem->MakeDataSlice(level);
        const unsigned long long words = (unsigned long long)((problem->graph->local_nodes + 31) >> 5);
        AbsorbDenseKernel<VertexId, SizeT, DataSlice, 256><<<util::FillGrid(words), 256, 0, stream>>>(
            problem->recv_map.ptr, problem->graph->part.Parts(), problem->visited_mask.ptr,
            problem->frontier_map[cur_out].ptr, problem->frontier_queues[in_sel ^ 1].ptr,
            problem->frontier_elements[in_sel ^ 1], problem->graph->row_offsets.ptr, Counters(), ds);
        GR_TRY(cudaGetLastError());
        in_sel ^= 1;
        ++enactor_stats.launches;
        return cudaSuccess;
    }
    // pull level: reads the (all-gathered) frontier map `cur`, writes this device's segment of map cur^1
    cudaError_t Pull(Problem *problem, VertexId level, int cur)
    {
        GR_TRY(ZeroCounters());
        DataSlice ds = problem->MakeDataSlice(level);
        typename Problem::Base::Graph &g = *problem->graph;
        oprtr::advance::pull::PullArgs<VertexId, SizeT> pa;
        pa.g = g.InverseView();
        pa.frontier_in  = problem->frontier_map[cur].ptr;
        pa.frontier_out = problem->frontier_map[cur ^ 1].ptr;
        pa.visited      = problem->visited_mask.ptr;
        pa.candidates   = problem->candidates.ptr;
        pa.head         = problem->head.ptr;
        pa.tickets      = nullptr;
        pa.fwd_row_offsets = g.symmetric ? nullptr : g.row_offsets.ptr;
        pa.out_queue = nullptr; pa.out_cap = 0;
        pa.slots = work_progress.d_slots.ptr; pa.queue_index = 0; pa.iteration = level;
        ++enactor_stats.launches;
        return oprtr::advance::LaunchPullKernel<PullPolicy, Problem, PullFunctor>(dev, ds, pa, stream);
    }
    cudaError_t MergeVisited(Problem *problem, int cur)
    {
        const unsigned long long words = (unsigned long long)((problem->graph->nodes + 31) >> 5);
        MergeVisitedKernel<<<util::FillGrid(words), 256, 0, stream>>>(problem->visited_mask.ptr,
                                                                      problem->frontier_map[cur].ptr, words);
        ++enactor_stats.launches;
        return util::GRError(cudaGetLastError(), "MergeVisitedKernel failed", __FILE__, __LINE__);
    }
    // local queue -> this device's segment of map `cur` (the whole map is cleared first)
    cudaError_t QueueToMap(Problem *problem, int cur, unsigned long long queue_len)
    {
        GR_TRY(problem->frontier_map[cur].Zero(stream));
        if (queue_len == 0) return cudaSuccess;
        DataSlice ds = problem->MakeDataSlice(0);
        QueueToBitmapKernel<VertexId, DataSlice><<<util::FillGrid(queue_len), 256, 0, stream>>>(
            problem->frontier_queues[in_sel].ptr, queue_len, problem->frontier_map[cur].ptr, ds);
        ++enactor_stats.launches;
        return util::GRError(cudaGetLastError(), "QueueToBitmapKernel failed", __FILE__, __LINE__);
    }
    // this device's segment of map `cur` -> local queue (length known to the caller from the counters)
    cudaError_t MapToQueue(Problem *problem, int cur)
    {
        DataSlice ds = problem->MakeDataSlice(0);
        oprtr::LevelCounters *scratch = work_progress.Slot(3);
        GR_TRY(cudaMemsetAsync(scratch, 0, sizeof(oprtr::LevelCounters), stream));
        const unsigned long long words = (unsigned long long)((problem->graph->local_nodes + 31) >> 5);
        BitmapToQueueKernel<VertexId, DataSlice, 256><<<util::FillGrid(words), 256, 0, stream>>>(
            problem->frontier_map[cur].ptr, problem->frontier_queues[in_sel].ptr,
            problem->frontier_elements[in_sel], scratch, ds);
        ++enactor_stats.launches;
        return util::GRError(cudaGetLastError(), "BitmapToQueueKernel failed", __FILE__, __LINE__);
    }
};

}  // namespace bfs
}  // namespace app
}  // namespace gunrock
