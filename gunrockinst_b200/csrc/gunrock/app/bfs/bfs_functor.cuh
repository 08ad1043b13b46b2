// bfs_functor.cuh -- the per-edge / per-vertex logic of BFS.
//
// Counterparts in the reference:
//   BFSFunctor          gunrock/app/bfs/bfs_functor.cuh:49-121
//   ReverseBFSFunctor   gunrock/app/dobfs/dobfs_functor.cuh:126-200
// Same four static device functions, same meaning of s_id (the source vertex when
// Problem::MARK_PREDECESSORS, else the iteration number).  What changes is the
// claim word: the reference wins a vertex with atomicCAS on its 4-byte label
// (or predecessor); here the winner is whoever flips the vertex's bit in the
// visited bitmask (one atomicOr on a word that is L2 resident), after which the
// label / predecessor are plain stores by a unique owner.
#pragma once
#include "bfs_problem.cuh"
#include "../../util/device_intrinsics.cuh"

namespace gunrock {
namespace app {
namespace bfs {

template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct BFSFunctor {
    typedef typename ProblemData::DataSlice DataSlice;

    // may `d_id` join the next frontier through this edge?
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        if (ProblemData::ENABLE_IDEMPOTENCE) return true;   // the filter decides (duplicates tolerated)
        return util::BitTestAndSet(problem->d_visited_mask, problem->BitPos(d_id));
    }

    // optional fifth member, used by the streaming advance: == active && CondEdge(...), with the claim issued
    // under a predicate so that the claims of a thread's entries overlap (util::AtomicOrIf)
    static __device__ __forceinline__ bool CondEdgeIf(
        bool active, VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        if (ProblemData::ENABLE_IDEMPOTENCE) return active;
        return util::BitTestAndSetIf(active, problem->d_visited_mask, problem->BitPos(d_id));
    }

    // the edge won `d_id`: record depth (and parent)
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        if (ProblemData::ENABLE_IDEMPOTENCE) return;
        const long long row = problem->LocalRow(d_id);
        util::StoreCG(problem->d_labels + row, (VertexId)(problem->level + 1));
        if (ProblemData::MARK_PREDECESSORS) util::StoreCG(problem->d_preds + row, s_id);
    }

    // drop holes left by failed edges
    static __device__ __forceinline__ bool CondFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0)
    {
        return node != -1;
    }

    // idempotent mode: the first filter pass that sees `node` unlabelled labels it
    // with `v` (the depth, or the predecessor when predecessors are marked --
    // reference filter/cta.cuh:229-231)
    static __device__ __forceinline__ void ApplyFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0)
    {
        if (ProblemData::ENABLE_IDEMPOTENCE) util::StoreCG(problem->d_labels + problem->LocalRow(node), (VertexId)v);
    }
};

// pull direction: `s_id` (parent, or iteration) reaches the unvisited `d_id`
template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct ReverseBFSFunctor {
    typedef typename ProblemData::DataSlice DataSlice;

    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        if (ProblemData::MARK_PREDECESSORS) util::StoreCG(problem->d_preds + problem->LocalRow(d_id), s_id);
        return true;
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        util::StoreCG(problem->d_labels + problem->LocalRow(d_id), (VertexId)(problem->level + 1));
    }
    static __device__ __forceinline__ bool CondFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0)
    {
        return node != -1;
    }
    static __device__ __forceinline__ void ApplyFilter(
        VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) {}
};

}  // namespace bfs
}  // namespace app
}  // namespace gunrock
