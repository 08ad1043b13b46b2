// sssp_functor.cuh -- per-edge relaxation and the priority score of SSSP.
//
// Counterparts in the reference: SSSPFunctor (gunrock/app/sssp/sssp_functor.cuh:36-112)
// and PQFunctor (:115-140).  Same contract; s_id is the source vertex, e_id the CSR entry
// (it indexes the weights).
#pragma once
#include "sssp_problem.cuh"

namespace gunrock {
namespace app {
namespace sssp {

template <typename VertexId, typename SizeT, typename ProblemData>
struct SSSPFunctor {
    typedef typename ProblemData::DataSlice DataSlice;

    // relax (s_id -> d_id); true when the distance of d_id improved AND d_id is not yet in
    // the frontier under construction (one queue entry per vertex and iteration; the
    // reference removes such duplicates afterwards with its MarkVisit lookup table,
    // priority_queue/kernel.cuh:116-140)
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        const unsigned label = problem->Dist(s_id);
        const unsigned weight = util::LoadStream(problem->d_weights + e_id);
        const unsigned new_weight = label + weight;          // unsigned, no guard: sssp_functor.cuh:60-62
        bool improved;
        if (ProblemData::MARK_PATHS) {
            // cheap pre-test on the distance half keeps most failing edges away from the atomic
            if (new_weight > (unsigned)(util::LoadCG(problem->d_dist_pred + d_id) >> 32)) return false;
            const unsigned long long word = ((unsigned long long)new_weight << 32) | (unsigned long long)(unsigned)s_id;
            const unsigned long long old = atomicMin(problem->d_dist_pred + d_id, word);
            improved = new_weight < (unsigned)(old >> 32);
        } else {
            if (new_weight >= util::LoadCG(problem->d_labels + d_id)) return false;
            improved = new_weight < atomicMin(problem->d_labels + d_id, new_weight);
        }
        return improved && util::BitTestAndSet(problem->d_mark, problem->BitPos(d_id));
    }

    // the parent travels inside the 64-bit word of CondEdge: nothing left to do
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0) {}

    static __device__ __forceinline__ bool CondFilter(
        VertexId node, const DataSlice *problem, unsigned v = 0, SizeT nid = 0) { return node != -1; }
    static __device__ __forceinline__ void ApplyFilter(
        VertexId node, const DataSlice *problem, unsigned v = 0, SizeT nid = 0) {}
};

template <typename VertexId, typename SizeT, typename ProblemData>
struct PQFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    // bucket of a vertex: distance / delta (reference: float division, :129-139)
    static __device__ __forceinline__ unsigned ComputePriorityScore(VertexId node_id, const DataSlice *problem)
    {
        return problem->Dist(node_id) / problem->delta;
    }
};

}  // namespace sssp
}  // namespace app
}  // namespace gunrock
