// bc_problem.cuh -- per-run device state of (Brandes) betweenness centrality.
//
// Counterpart of the reference's BCProblem (gunrock/app/bc/bc_problem.cuh:33-485):
//   d_labels     int32 [n]   depth from the current source, -1 = unreached   (:351)
//   d_sigmas     fp32  [n]   number of shortest paths from the source        (:421, source = 1: :464)
//   d_deltas     fp32  [n]   dependency of the source on the vertex          (:392)
//   d_bc_values  fp32  [n]   accumulated over sources, NOT cleared by Reset  (:336)
//   d_ebc_values fp32  [m]   allocated and zeroed by the reference but never written (its
//                            accumulation is commented out, bc_functor.cuh:207,313); the C entry
//                            point therefore returns zeros for it
// The reference re-discovers each level's frontier in the backward pass by filtering the
// whole vertex set on its label (BackwardFunctor::CondFilter); here the forward pass appends
// every level to ONE queue of n ids (a vertex is discovered once), so the backward pass walks
// the levels it already has, deepest first.
#pragma once
#include <vector>
#include "../problem_base.cuh"

namespace gunrock {
namespace app {
namespace bc {

template <typename _VertexId, typename _SizeT, typename _Value>
struct BCProblem : ProblemBase<_VertexId, _SizeT> {
    typedef _VertexId VertexId;
    typedef _SizeT    SizeT;
    typedef _Value    Value;
    typedef ProblemBase<VertexId, SizeT> Base;
    static const bool MARK_PREDECESSORS  = true;    // functors need the real source vertex (reference: bc_app.cu:65)
    static const bool ENABLE_IDEMPOTENCE = false;

    struct DataSlice {
        VertexId *d_labels;
        Value    *d_sigmas, *d_deltas, *d_bc_values;
        VertexId  src;
        VertexId  level;            // depth of the frontier being expanded
        long long nodes;
    };

    util::DeviceArray<VertexId> labels, level_queue;
    util::DeviceArray<Value>    sigmas, deltas, bc_values;
    VertexId src = 0;
    unsigned long long src_degree = 0;

    cudaError_t Init(typename Base::Graph *g, cudaStream_t s)
    {
        GR_TRY(Base::InitBase(g, s));
        const size_t n = (size_t)g->local_nodes;
        GR_TRY(labels.Allocate(n));
        GR_TRY(level_queue.Allocate(n + 1));
        GR_TRY(sigmas.Allocate(n));
        GR_TRY(deltas.Allocate(n));
        GR_TRY(bc_values.Allocate(n));
        return bc_values.Zero(s);                    // accumulates across Reset/Enact pairs
    }

    DataSlice MakeDataSlice(VertexId level) const
    {
        DataSlice d;
        d.d_labels = labels.ptr; d.d_sigmas = sigmas.ptr; d.d_deltas = deltas.ptr; d.d_bc_values = bc_values.ptr;
        d.src = src; d.level = level; d.nodes = this->graph->nodes;
        return d;
    }

    // reference: BCProblem::Reset (bc_problem.cuh:324-470)
    cudaError_t Reset(VertexId source)
    {
        cudaStream_t s = this->stream;
        src = source; src_degree = 0;
        const long long n = this->graph->local_nodes;
        if (n) GR_TRY(cudaMemsetAsync(labels.ptr, 0xff, sizeof(VertexId) * (size_t)n, s));
        GR_TRY(sigmas.Zero(s));
        GR_TRY(deltas.Zero(s));
        if (source < 0 || (long long)source >= this->graph->nodes) return cudaSuccess;
        const VertexId zero = 0;
        const Value one = (Value)1;
        GR_TRY(cudaMemcpyAsync(labels.ptr + source, &zero, sizeof(VertexId), cudaMemcpyHostToDevice, s));
        GR_TRY(cudaMemcpyAsync(sigmas.ptr + source, &one, sizeof(Value), cudaMemcpyHostToDevice, s));
        GR_TRY(cudaMemcpyAsync(level_queue.ptr, &source, sizeof(VertexId), cudaMemcpyHostToDevice, s));
        SizeT ends[2];
        GR_TRY(cudaMemcpyAsync(ends, this->graph->row_offsets.ptr + source, 2 * sizeof(SizeT), cudaMemcpyDeviceToHost, s));
        GR_TRY(cudaStreamSynchronize(s));
        src_degree = (unsigned long long)(ends[1] - ends[0]);
        return cudaSuccess;
    }

    // reference: BCProblem::Extract (bc_problem.cuh:152-190)
    cudaError_t Extract(Value *h_sigmas, Value *h_bc_values, VertexId *h_labels = nullptr)
    {
        cudaStream_t s = this->stream;
        const size_t n = (size_t)this->graph->local_nodes;
        if (h_sigmas) GR_TRY(sigmas.ToHost(h_sigmas, n, s));
        if (h_bc_values) GR_TRY(bc_values.ToHost(h_bc_values, n, s));
        if (h_labels) GR_TRY(labels.ToHost(h_labels, n, s));
        return cudaStreamSynchronize(s);
    }
};

}  // namespace bc
}  // namespace app
}  // namespace gunrock
