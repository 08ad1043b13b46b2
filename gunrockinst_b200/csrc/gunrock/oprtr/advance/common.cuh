// common.cuh -- pieces shared by the push-direction advance kernels (LB and TWC):
// the functor call convention and the CTA-level output stager.
//
// Functor contract (unchanged from the reference, SURVEY.md section 8 a11;
// e.g. gunrock/app/bfs/bfs_functor.cuh:49-121):
//   static bool CondEdge (VertexId s_id, VertexId d_id, const DataSlice*, SizeT e_id, SizeT e_id_in)
//   static void ApplyEdge(VertexId s_id, VertexId d_id, const DataSlice*, SizeT e_id, SizeT e_id_in)
// where s_id is the source vertex only when Problem::MARK_PREDECESSORS, else the
// iteration number (oprtr/edge_map_partitioned/kernel.cuh:401-403), e_id is the
// CSR entry index and e_id_in the position of the source in the input frontier.
// A failed CondEdge produces no output (OUT_COMPACT) or a -1 (OUT_TOMBSTONE).
//
// Output staging replaces the reference's "one slot per edge + raking scan +
// separate filter pass" (edge_map_forward/cta.cuh:743-756, filter/cta.cuh:
// 467-544): winners are packed into shared memory with one ballot per warp
// step and leave for HBM as coalesced runs, one global atomic per flush.
#pragma once
#include "kernel_policy.cuh"
#include "../../util/device_intrinsics.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {

template <typename Policy>
struct StagerStorage {
    typename Policy::VertexId stage[Policy::STAGE_CAP];
    unsigned                  count;
    unsigned long long        flush_base;
};

template <typename Policy>
struct Stager {
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    StagerStorage<Policy> &sm;
    const Args<VertexId, SizeT> &args;
    LevelCounters *slot_out;
    unsigned long long out_degree;      // per-thread partial (TRACK_OUT_DEGREE)

    __device__ __forceinline__ Stager(StagerStorage<Policy> &s, const Args<VertexId, SizeT> &a, LevelCounters *so)
        : sm(s), args(a), slot_out(so), out_degree(0) {}

    // call once, before the first Push, from all threads; followed by a __syncthreads by the caller
    __device__ __forceinline__ void Init() { if (threadIdx.x == 0) sm.count = 0; }

    // warp-convergent: every lane of the warp calls it, `ok` selects the lanes that emit
    __device__ __forceinline__ void Push(bool ok, VertexId v)
    {
        const unsigned m = __ballot_sync(util::kFullMask, ok);
        if (m) {
            const int lane = util::LaneId(), leader = __ffs(m) - 1;
            unsigned b = 0;
            if (lane == leader) b = atomicAdd(&sm.count, (unsigned)__popc(m));
            b = __shfl_sync(util::kFullMask, b, leader);
            if (ok) sm.stage[b + __popc(m & ((1u << lane) - 1))] = v;
        }
    }

    // CTA-convergent: write everything staged so far to the output frontier
    __device__ __forceinline__ void Flush()
    {
        __syncthreads();
        const unsigned cnt = sm.count;      // same value in every thread
        __syncthreads();                    // ... and nobody pushes again before all have read it
        if (cnt != 0) {
            if (threadIdx.x == 0) {
                sm.flush_base = atomicAdd(&slot_out->out_len, (unsigned long long)cnt);
                sm.count = 0;               // every thread already holds cnt
            }
            __syncthreads();
            const unsigned long long fb = sm.flush_base;
            for (unsigned k = threadIdx.x; k < cnt; k += Policy::THREADS) {
                VertexId v = sm.stage[k];
                if (fb + k < args.out_cap) args.out_queue[fb + k] = v;
                else slot_out->overflow = 1;
                if constexpr (Policy::TRACK_OUT_DEGREE) {
                    if (OutputIsEdges(args.type)) v = args.g.col_indices[v];       // degree of the vertex the edge points to
                    if (args.g.part.Mine(v)) {
                        long long row = args.g.part.Local(v);
                        out_degree += (unsigned long long)(args.g.row_offsets[row + 1] - args.g.row_offsets[row]);
                    }
                }
            }
            __syncthreads();                // stage[] may be overwritten from here on
        }
    }

    // as FlushIfNeeded, for a caller that has just passed a __syncthreads() and does not Push
    // again before its next one (the count is then stable without two more barriers)
    __device__ __forceinline__ void FlushIfNeededSynced(unsigned incoming)
    {
        if (sm.count + incoming > (unsigned)Policy::STAGE_CAP) Flush();
    }

    // CTA-convergent: flush only if another `incoming` entries might not fit
    __device__ __forceinline__ void FlushIfNeeded(unsigned incoming)
    {
        __syncthreads();
        const unsigned cnt = sm.count;
        __syncthreads();
        if (cnt + incoming > (unsigned)Policy::STAGE_CAP) Flush();
    }
};

// evaluate the functor pair on one edge
template <typename Policy, typename Problem, typename Functor>
__device__ __forceinline__ bool VisitEdge(
    const typename Problem::DataSlice *ds, typename Policy::VertexId s_id, typename Policy::VertexId d_id,
    typename Policy::SizeT e_id, typename Policy::SizeT e_in)
{
    if (Functor::CondEdge(s_id, d_id, ds, e_id, e_in)) {
        Functor::ApplyEdge(s_id, d_id, ds, e_id, e_in);
        return true;
    }
    return false;
}

// active && CondEdge(...): through the functor's optional predicated form (CondEdgeIf) when it has one
template <typename Functor, typename DataSlice, typename VertexId, typename SizeT, typename = void>
struct CondEdgeIfImpl {
    static __device__ __forceinline__ bool Call(bool active, VertexId s, VertexId d, const DataSlice *ds, SizeT e, SizeT e_in)
    {
        return active && Functor::CondEdge(s, d, ds, e, e_in);
    }
};
template <typename Functor, typename DataSlice, typename VertexId, typename SizeT>
struct CondEdgeIfImpl<Functor, DataSlice, VertexId, SizeT,
                      decltype((void)Functor::CondEdgeIf(true, VertexId(), VertexId(), (const DataSlice *)nullptr, SizeT(), SizeT()), void())> {
    static __device__ __forceinline__ bool Call(bool active, VertexId s, VertexId d, const DataSlice *ds, SizeT e, SizeT e_in)
    {
        return Functor::CondEdgeIf(active, s, d, ds, e, e_in);
    }
};

// CTA-wide sum + one atomic into a 64-bit counter (CTA-convergent)
template <int THREADS>
__device__ __forceinline__ void BlockAccumulate(unsigned long long v, unsigned long long *scratch, unsigned long long *dst)
{
    unsigned long long total;
    (void)util::BlockExclusiveSum<THREADS>(v, scratch, total);
    if (threadIdx.x == 0 && total) atomicAdd(dst, total);
}

}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
