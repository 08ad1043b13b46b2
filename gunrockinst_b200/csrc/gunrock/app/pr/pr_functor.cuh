// pr_functor.cuh -- per-edge and per-vertex logic of PageRank.
//
// Counterparts in the reference (gunrock/app/pr/pr_functor.cuh): PRFunctor (:32-100) and
// RemoveZeroDegreeNodeFunctor (:113-175).  Same contract and the same arithmetic, in fp64:
//   CondEdge(s, d)   both end points still have out-edges
//   ApplyEdge(s, d)  rank_next[d] += rank_curr[s] / degree[s]      (one atomicAdd per edge)
//   CondFilter(v)    rank_next[v] = delta * rank_next[v] + (1 - delta) * [src == v or src == -1];
//                    survives (is counted as still moving) when |next - curr| > threshold
// ApplyFilter caches next / degree for the following iteration's edges.
#pragma once
#include "pr_problem.cuh"
#include "../../util/device_intrinsics.cuh"

namespace gunrock {
namespace app {
namespace pr {

template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct PRFunctor {
    typedef typename ProblemData::DataSlice DataSlice;

    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        // d_contrib[s] is zero when s was pruned, so only the destination needs a look
        return util::LoadCA(problem->d_degrees + d_id) > 0;
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        atomicAdd(problem->d_rank_next + d_id, util::LoadCA(problem->d_contrib + s_id));
    }
    static __device__ __forceinline__ bool CondFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0)
    {
        const Value delta = problem->delta;
        const Value base = (problem->src_node == node || problem->src_node == -1) ? 1.0 : 0.0;
        const Value next = __dadd_rn(__dmul_rn(delta, problem->d_rank_next[node]), __dmul_rn(1.0 - delta, base));
        problem->d_rank_next[node] = next;
        problem->d_contrib[node] = next / (Value)problem->d_degrees[node];
        return fabs(next - problem->d_rank_curr[node]) > problem->threshold;
    }
    static __device__ __forceinline__ void ApplyFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) {}
};

template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct RemoveZeroDegreeNodeFunctor {
    typedef typename ProblemData::DataSlice DataSlice;

    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        return util::LoadCA(problem->d_degrees + d_id) == 0;
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        atomicAdd(problem->d_degrees_pong + s_id, -1);
    }
    static __device__ __forceinline__ bool CondFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0)
    {
        const int d = problem->d_degrees[node];
        if (d == 0) problem->d_degrees_pong[node] = -1;
        return d > 0;
    }
    static __device__ __forceinline__ void ApplyFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) {}
};

}  // namespace pr
}  // namespace app
}  // namespace gunrock
