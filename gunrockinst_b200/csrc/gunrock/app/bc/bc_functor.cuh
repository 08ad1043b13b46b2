// bc_functor.cuh -- the per-edge logic of betweenness centrality.
//
// Counterparts in the reference: ForwardFunctor (gunrock/app/bc/bc_functor.cuh:36-119) and
// BackwardFunctor (:130-222).  Same contract (s_id is the real source vertex).
//   forward   (s at depth L, edge s->d): the first edge to reach d claims it for depth L+1 and
//             puts it in the next frontier; EVERY edge into a depth L+1 vertex adds sigma[s]
//             to sigma[d] (the reference does the add in ApplyEdge for the winner and in the
//             failing branch of CondEdge for the others)
//   backward  (s at depth L, edge s->d with d at depth L+1):
//             delta[s] += sigma[s] / sigma[d] * (1 + delta[d]), and the same amount into
//             bc[s] unless s is the source (:195-200)
#pragma once
#include "bc_problem.cuh"
#include "../../util/device_intrinsics.cuh"

namespace gunrock {
namespace app {
namespace bc {

template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct ForwardFunctor {
    typedef typename ProblemData::DataSlice DataSlice;

    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        const VertexId next = problem->level + 1;
        VertexId seen = util::LoadCG(problem->d_labels + d_id);
        bool won = false;
        if (seen == -1) {
            seen = atomicCAS(problem->d_labels + d_id, (VertexId)-1, next);
            won = seen == -1;
            if (won) seen = next;
        }
        // sigma[s_id] is final: s_id was completed by the previous level
        if (seen == next) atomicAdd(problem->d_sigmas + d_id, util::LoadCG(problem->d_sigmas + s_id));
        return won;
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0) {}
    static __device__ __forceinline__ bool CondFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) {}
};

template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct BackwardFunctor {
    typedef typename ProblemData::DataSlice DataSlice;

    // is d_id a successor of s_id on a shortest path?  (problem->level is the depth of s_id)
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        return util::LoadCG(problem->d_labels + d_id) == problem->level + 1;
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        if (s_id == problem->src) return;
        const Value from_sigma = util::LoadCG(problem->d_sigmas + s_id);
        const Value to_sigma   = util::LoadCG(problem->d_sigmas + d_id);
        const Value to_delta   = util::LoadCG(problem->d_deltas + d_id);     // final: deeper levels are done
        const Value result = from_sigma / to_sigma * ((Value)1 + to_delta);
        atomicAdd(problem->d_deltas + s_id, result);
        atomicAdd(problem->d_bc_values + s_id, result);
    }
    static __device__ __forceinline__ bool CondFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) {}
};

}  // namespace bc
}  // namespace app
}  // namespace gunrock
