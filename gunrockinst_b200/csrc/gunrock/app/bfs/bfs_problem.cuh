// bfs_problem.cuh -- per-run device state of breadth-first search.
//
// Counterpart of the reference's BFSProblem (gunrock/app/bfs/bfs_problem.cuh:
// 33-370) and DOBFSProblem (app/dobfs/dobfs_problem.cuh:35-420) in one type:
//   d_labels        int32 [local_nodes]  depth, -1 = unreached        (bfs_problem.cuh:298)
//   d_preds         int32 [local_nodes]  parent, -2 = unreached, source -1 (:307,342-348)
//   d_visited_mask  1 bit / vertex, GLOBAL (replicated cull hints on a partitioned
//                   run; the owner's words are authoritative).  The reference
//                   keeps a byte-addressed mask only for idempotent runs (:317);
//                   here it is also the unique-claim word of the normal mode.
//   d_frontier_map  2 x 1 bit / vertex, GLOBAL: in / out frontier of the pull
//                   direction (reference: two bool-BYTE arrays,
//                   dobfs_problem.cuh:63-64)
//   d_candidates    1 bit / local vertex: has at least one in-edge
// Bit position of vertex v: Owner(v) * local_nodes + Local(v) ("owner-major"), so
// a device's bits are one contiguous run that collectives can move in place.
#pragma once
#include "../problem_base.cuh"

namespace gunrock {
namespace app {
namespace bfs {

template <typename VertexId>
__global__ void MarkCandidatesKernel(const void *offsets, int offset_bytes, long long local_nodes, unsigned *cand)
{
    const long long words = (local_nodes + 31) >> 5;
    const long long stride = (long long)gridDim.x * blockDim.x / 32;
    const int lane = threadIdx.x & 31;
    for (long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / 32; w < words; w += stride) {
        const long long row = (w << 5) + lane;
        bool has = false;
        if (row < local_nodes) {
            if (offset_bytes == 8) has = ((const long long *)offsets)[row + 1] > ((const long long *)offsets)[row];
            else                   has = ((const int *)offsets)[row + 1] > ((const int *)offsets)[row];
        }
        const unsigned m = __ballot_sync(0xffffffffu, has);
        if (lane == 0) cand[w] = m;
    }
}

// head[row] = first entry of the row's (inverse) list, -1 if the row is empty
template <typename VertexId, typename SizeT>
__global__ void FillHeadKernel(const SizeT *offsets, const VertexId *cols, long long local_nodes, VertexId *head)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x; row < local_nodes; row += stride) {
        const SizeT b = offsets[row], e = offsets[row + 1];
        head[row] = e > b ? cols[b] : (VertexId)-1;
    }
}

template <typename _VertexId, typename _SizeT, typename _Value, bool _MARK_PREDECESSORS, bool _ENABLE_IDEMPOTENCE>
struct BFSProblem : ProblemBase<_VertexId, _SizeT> {
    typedef _VertexId VertexId;
    typedef _SizeT    SizeT;
    typedef _Value    Value;
    typedef ProblemBase<VertexId, SizeT> Base;
    static const bool MARK_PREDECESSORS  = _MARK_PREDECESSORS;
    static const bool ENABLE_IDEMPOTENCE = _ENABLE_IDEMPOTENCE;

    // what the functors see (copied into the kernel parameter bank at every launch)
    struct DataSlice {
        VertexId  *d_labels;
        VertexId  *d_preds;
        unsigned  *d_visited_mask;
        VertexId   level;           // depth of the frontier being expanded
        long long  nodes, local_nodes;
        graph::Partition part;
        // partitioned push: discoveries (of any owner) are set in this GLOBAL bitmap, whose
        // owner-major segments are then exchanged all-to-all (NULL on a single device)
        unsigned  *d_next_map;

        __host__ __device__ __forceinline__ unsigned long long BitPos(VertexId v) const
        {
            if (part.log_parts == 0) return (unsigned long long)(unsigned)v;
            return (unsigned long long)part.Owner(v) * (unsigned long long)local_nodes + (unsigned long long)part.Local(v);
        }
        __host__ __device__ __forceinline__ long long LocalRow(VertexId v) const { return part.Local(v); }
    };

    util::DeviceArray<VertexId> labels, preds;
    util::DeviceArray<unsigned> visited_mask, frontier_map[2], candidates;
    util::DeviceArray<VertexId> head;                   // pull direction: first in-neighbour of every row
    util::DeviceArray<unsigned> next_map, recv_map;     // partitioned runs only (PrepareExchange)
    bool candidates_valid = false;
    // the graph is the degree-ordered twin of the caller's graph: caller's id -> id here (device);
    // applied by the persistent kernel to the sources of a batch (single runs translate on the host)
    const VertexId *src_map = nullptr;
    // ... and what every search ends with in that case: labels (and predecessors) gathered back into the
    // caller's ids, out_labels[v] = labels[src_map[v]] (predecessor VALUES translated through dst_map)
    const VertexId *dst_map = nullptr;
    util::DeviceArray<VertexId> out_labels, out_preds;
    cudaError_t SetTwinMaps(const VertexId *to_new, const VertexId *to_old)
    {
        src_map = to_new; dst_map = to_old;
        if (!to_new) return cudaSuccess;
        GR_TRY(out_labels.Allocate((size_t)this->graph->nodes + 1));
        if (MARK_PREDECESSORS) GR_TRY(out_preds.Allocate((size_t)this->graph->nodes + 1));
        return cudaSuccess;
    }
    unsigned long long src_degree = 0;
    VertexId src = 0;

    size_t MaskWords() const { return (size_t)((this->graph->nodes + 31) / 32) + 1; }

    cudaError_t AllocateState()
    {
        size_t ln = (size_t)this->graph->local_nodes;
        GR_TRY(labels.Allocate(ln));
        if (MARK_PREDECESSORS) GR_TRY(preds.Allocate(ln));
        GR_TRY(visited_mask.Allocate(MaskWords()));
        candidates_valid = false;
        return cudaSuccess;
    }
    // resident graph
    cudaError_t Init(typename Base::Graph *g, cudaStream_t s) { GR_TRY(Base::InitBase(g, s)); return AllocateState(); }
    // host CSR (C API path)
    template <typename HostSizeT>
    cudaError_t Init(long long n, long long m, const HostSizeT *h_off, const VertexId *h_col, int device, cudaStream_t s)
    {
        GR_TRY(Base::InitBase(n, m, h_off, h_col, (const unsigned *)nullptr, device, s));
        return AllocateState();
    }

    // state for the pull direction (allocated on first use)
    cudaError_t PreparePull()
    {
        GR_TRY(frontier_map[0].Allocate(MaskWords()));
        GR_TRY(frontier_map[1].Allocate(MaskWords()));
        if (!candidates_valid) {
            size_t words = (size_t)((this->graph->local_nodes + 31) / 32) + 1;
            GR_TRY(candidates.Allocate(words));
            auto inv = this->graph->InverseView();
            MarkCandidatesKernel<VertexId><<<util::FillGrid(words * 32), 256, 0, this->stream>>>(
                (const void *)inv.row_offsets, (int)sizeof(SizeT), this->graph->local_nodes, candidates.ptr);
            GR_TRY(cudaGetLastError());
            GR_TRY(head.Allocate((size_t)this->graph->local_nodes + 1));
            FillHeadKernel<VertexId, SizeT><<<util::FillGrid((unsigned long long)this->graph->local_nodes + 1), 256, 0,
                                               this->stream>>>(inv.row_offsets, inv.col_indices,
                                                               this->graph->local_nodes, head.ptr);
            GR_TRY(cudaGetLastError());
            candidates_valid = true;
        }
        return cudaSuccess;
    }

    // state for a 1D-partitioned run: the dense discovery bitmap a push level fills and the
    // buffer its segments are received into (parts x own-segment words)
    cudaError_t PrepareExchange()
    {
        GR_TRY(next_map.Allocate(MaskWords()));
        return recv_map.Allocate(MaskWords());
    }

    DataSlice MakeDataSlice(VertexId level) const
    {
        DataSlice d;
        d.d_labels = labels.ptr; d.d_preds = preds.ptr; d.d_visited_mask = visited_mask.ptr;
        d.level = level; d.nodes = this->graph->nodes; d.local_nodes = this->graph->local_nodes;
        d.part = this->graph->part;
        d.d_next_map = next_map.size ? next_map.ptr : nullptr;
        return d;
    }

    // reference: BFSProblem::Reset (bfs_problem.cuh:272-357): labels <- -1, preds <- -2,
    // mask <- 0, queue[0] = {src}, labels[src] = 0, preds[src] = -1
    // on_device: only size the queues and note the source; the persistent BFS kernel performs the
    // fills below itself (bfs_persistent.cuh), which spares ~8 API calls and a host sync per search
    cudaError_t Reset(VertexId source, FrontierType ftype, double queue_sizing, bool on_device = false)
    {
        GR_TRY(Base::ResetBase(ftype, queue_sizing));
        src = source;
        if (on_device) { src_degree = 0; return cudaSuccess; }
        return FillState(source);
    }

    // the device work of Reset (what the persistent kernel otherwise does itself)
    cudaError_t FillState(VertexId source)
    {
        cudaStream_t s = this->stream;
        const long long ln = this->graph->local_nodes;
        if (ln) GR_TRY(cudaMemsetAsync(labels.ptr, 0xff, sizeof(VertexId) * (size_t)ln, s));
        if (MARK_PREDECESSORS) GR_TRY(util::Fill(preds.ptr, (VertexId)-2, (unsigned long long)ln, s));
        GR_TRY(visited_mask.Zero(s));
        src_degree = 0;
        if (source < 0 || (long long)source >= this->graph->nodes) return cudaSuccess;   // empty search
        const graph::Partition &pt = this->graph->part;
        DataSlice d = MakeDataSlice(0);
        // every device marks the source in its replicated mask; the owner seeds labels + queue
        unsigned long long pos = d.BitPos(source);
        unsigned word = 1u << (pos & 31);
        GR_TRY(cudaMemcpyAsync(visited_mask.ptr + (pos >> 5), &word, sizeof(unsigned), cudaMemcpyHostToDevice, s));
        if (pt.Mine(source)) {
            const long long row = pt.Local(source);
            VertexId zero = 0, minus1 = -1;
            GR_TRY(cudaMemcpyAsync(labels.ptr + row, &zero, sizeof(VertexId), cudaMemcpyHostToDevice, s));
            if (MARK_PREDECESSORS)
                GR_TRY(cudaMemcpyAsync(preds.ptr + row, &minus1, sizeof(VertexId), cudaMemcpyHostToDevice, s));
            GR_TRY(cudaMemcpyAsync(this->frontier_queues[0].ptr, &source, sizeof(VertexId), cudaMemcpyHostToDevice, s));
            SizeT ends[2];
            GR_TRY(cudaMemcpyAsync(ends, this->graph->row_offsets.ptr + row, 2 * sizeof(SizeT), cudaMemcpyDeviceToHost, s));
            GR_TRY(cudaStreamSynchronize(s));
            src_degree = (unsigned long long)(ends[1] - ends[0]);
        } else {
            GR_TRY(cudaStreamSynchronize(s));       // the host words above must outlive the copies
        }
        return cudaSuccess;
    }

    // reference: BFSProblem::Extract (bfs_problem.cuh:144-177)
    cudaError_t Extract(VertexId *h_labels, VertexId *h_preds)
    {
        cudaStream_t s = this->stream;
        GR_TRY(labels.ToHost(h_labels, (size_t)this->graph->local_nodes, s));
        if (MARK_PREDECESSORS && h_preds) GR_TRY(preds.ToHost(h_preds, (size_t)this->graph->local_nodes, s));
        return cudaStreamSynchronize(s);
    }
};

}  // namespace bfs
}  // namespace app
}  // namespace gunrock
