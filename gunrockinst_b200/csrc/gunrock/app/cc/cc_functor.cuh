// cc_functor.cuh -- hooking and pointer jumping of connected components.
//
// The reference (Soman et al.) alternates bulk passes, each one filter launch with its own
// functor and a host-side flag round trip: HookInit, HookMax (HookMin is commented out),
// PtrJump, PtrJumpMask, PtrJumpUnmask, UpdateMask (gunrock/app/cc/cc_functor.cuh:19-410,
// cc_enactor.cuh:443-860) -- on the order of 10^2 launches for an R-MAT graph.  All of its
// hooks link the LARGER root under the SMALLER one, so labels converge to the minimum
// vertex id of each component.
//
// Here the same hook rule is applied asynchronously: an edge keeps chasing the two roots
// (with path halving) until they coincide or its compare-and-swap links the larger root
// under the smaller one, so ONE pass over the edges plus ONE flattening pass over the
// vertices reach the same fixed point -- identical labels, no flags, no host round trips.
// Parent pointers only ever move to ancestors and ancestors are always smaller ids, so the
// racing plain stores of path halving are benign.
#pragma once
#include "cc_problem.cuh"
#include "../../util/device_intrinsics.cuh"

namespace gunrock {
namespace app {
namespace cc {

template <typename VertexId>
__device__ __forceinline__ VertexId FindRoot(VertexId *parent, VertexId x)
{
    VertexId p = util::LoadCG(parent + x);
    while (p != x) {
        const VertexId gp = util::LoadCG(parent + p);
        if (gp != p) util::StoreCG(parent + x, gp);     // path halving
        x = p; p = gp;
    }
    return x;
}

// counterpart of HookInitFunctor / HookMaxFunctor (cc_functor.cuh:65-103,184-250), used
// through the ADVANCE operator: s_id, d_id are the end points of CSR entry e_id
template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct HookFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    static __device__ __forceinline__ bool CondEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0)
    {
        if (s_id == d_id || (problem->skip_upper && d_id > s_id)) return false;
        VertexId *parent = problem->d_component_ids;
        VertexId u = s_id, v = d_id;
        for (;;) {
            u = FindRoot(parent, u);
            v = FindRoot(parent, v);
            if (u == v) break;
            const VertexId hi = u > v ? u : v, lo = u > v ? v : u;
            const VertexId old = atomicCAS(parent + hi, hi, lo);
            if (old == hi) break;               // linked larger root -> smaller root
            u = old; v = lo;                    // `hi` stopped being a root meanwhile: chase again
        }
        return false;                           // nothing joins a frontier
    }
    static __device__ __forceinline__ void ApplyEdge(
        VertexId s_id, VertexId d_id, const DataSlice *problem, SizeT e_id = 0, SizeT e_id_in = 0) {}
    static __device__ __forceinline__ bool CondFilter(VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) { return true; }
    static __device__ __forceinline__ void ApplyFilter(VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0) {}
};

// counterpart of PtrJumpFunctor (cc_functor.cuh:116-170), used through the FILTER operator
// in apply-only mode: point every vertex at its root and count the roots
template <typename VertexId, typename SizeT, typename Value, typename ProblemData>
struct PtrJumpFunctor {
    typedef typename ProblemData::DataSlice DataSlice;
    static __device__ __forceinline__ bool CondFilter(VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0)
    {
        return true;
    }
    static __device__ __forceinline__ void ApplyFilter(VertexId node, const DataSlice *problem, Value v = 0, SizeT nid = 0)
    {
        VertexId *parent = problem->d_component_ids;
        VertexId r = util::LoadCG(parent + node);
        if (r != node) {
            VertexId p = util::LoadCG(parent + r);
            while (p != r) { r = p; p = util::LoadCG(parent + r); }
            util::StoreCG(parent + node, r);
        } else {
            atomicAdd(problem->d_num_components, 1ull);
        }
    }
    static __device__ __forceinline__ bool CondEdge(VertexId s, VertexId d, const DataSlice *problem, SizeT e = 0, SizeT ei = 0) { return false; }
    static __device__ __forceinline__ void ApplyEdge(VertexId s, VertexId d, const DataSlice *problem, SizeT e = 0, SizeT ei = 0) {}
};

}  // namespace cc
}  // namespace app
}  // namespace gunrock
