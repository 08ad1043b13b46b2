// reduce.cuh -- neighbour-reduce advance: one value per frontier item, reduced over the edges of
// that item the functor accepts.
//
// Counterpart of the reference's advance with R_TYPE / R_OP set (gunrock/oprtr/advance/
// kernel.cuh:731-761, oprtr/edge_map_partitioned/kernel.cuh:417-470): there the LB kernels store
// `d_value_to_reduce[u]` (VERTEX) or `[e]` (EDGE) -- or the operator's identity for an edge whose
// CondEdge failed -- into a per-edge scratch frontier, and a separate moderngpu SegReduceCsr pass
// reduces each item's segment.  Here both steps are one kernel with no per-edge scratch: a warp
// walks an item's list in 128-byte steps, folds accepted values in registers and finishes with a
// shuffle reduction; an item with no accepted edge yields the identity (PLUS / BIT_OR / BIT_XOR 0,
// MULTIPLIES 1, MAXIMUM INT_MIN, MINIMUM INT_MAX, BIT_AND ~0, as the reference).
// Frontier types (V2V / V2E / E2V / E2E) and the optional compacted output queue behave as in
// the other push kernels.  Algorithmic bytes: 4 + 2*sizeof(SizeT) per item, 4 + sizeof(Value)
// per inspected edge, sizeof(Value) per item written.
#pragma once
#include <climits>
#include <type_traits>
#include "common.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {

// names and order follow the reference (oprtr/advance/kernel_policy.cuh:62-81)
enum REDUCE_OP { NONE, PLUS, MINUS, MULTIPLIES, MODULUS, BIT_OR, BIT_AND, BIT_XOR, MAXIMUM, MINIMUM };
enum REDUCE_TYPE { EMPTY, VERTEX, EDGE };

template <typename Value>
struct ReduceArgs {
    const Value *values;        // indexed by destination vertex (VERTEX) or CSR entry (EDGE)
    Value       *reduced;       // [in_len] one result per input item
    REDUCE_OP    op;
    REDUCE_TYPE  rtype;
};

template <typename Value>
__host__ __device__ __forceinline__ Value ReduceIdentity(REDUCE_OP op)
{
    switch (op) {
        case MULTIPLIES: return (Value)1;
        case MAXIMUM:    return (Value)INT_MIN;
        case MINIMUM:    return (Value)INT_MAX;
        case BIT_AND:    return (Value)0xffffffff;
        default:         return (Value)0;
    }
}
template <typename Value>
__host__ __device__ __forceinline__ Value ReduceCombine(REDUCE_OP op, Value a, Value b)
{
    switch (op) {
        case MULTIPLIES: return a * b;
        case MAXIMUM:    return a > b ? a : b;
        case MINIMUM:    return a < b ? a : b;
        default: break;
    }
    if constexpr (std::is_integral<Value>::value) {
        if (op == BIT_OR)  return a | b;
        if (op == BIT_AND) return a & b;
        if (op == BIT_XOR) return a ^ b;
    }
    return a + b;               // PLUS, and the reference's default for the operators it never wired up
}

template <typename Policy, typename Problem, typename Functor, typename Value>
__global__ void __launch_bounds__(Policy::THREADS, Policy::MIN_CTAS)
NeighborReduceKernel(const __grid_constant__ typename Problem::DataSlice ds,
                     const __grid_constant__ Args<typename Policy::VertexId, typename Policy::SizeT> args,
                     const __grid_constant__ ReduceArgs<Value> red)
{
    typedef typename Policy::VertexId VertexId;
    typedef typename Policy::SizeT    SizeT;
    constexpr int THREADS = Policy::THREADS;
    __shared__ unsigned long long s_red[THREADS / 32 + 1];
    const int lane = util::LaneId();
    const unsigned long long in_len = args.d_in_len ? *args.d_in_len : args.in_len;
    LevelCounters *slot_out = args.slots + ((args.queue_index + 1) & 3);
    if (blockIdx.x == 0 && threadIdx.x == 0) ClearFutureSlot(args.slots, args.queue_index);
    const unsigned long long warps = (unsigned long long)gridDim.x * (THREADS / 32);
    unsigned long long my_edges = 0;

    for (unsigned long long i = (unsigned long long)blockIdx.x * (THREADS / 32) + util::WarpId(); i < in_len; i += warps) {
        VertexId v = args.in_queue ? util::LoadCG(args.in_queue + i) : (VertexId)args.g.part.Global((long long)i);
        if (InputIsEdges(args.type) && args.in_queue)
            v = (v >= 0 && (long long)v < args.g.edges) ? args.g.col_indices[v] : (VertexId)-1;
        long long begin = 0, end = 0;
        if (v >= 0 && (long long)v < args.g.nodes) {
            const long long row = args.g.part.Local(v);
            begin = (long long)args.g.row_offsets[row];
            end   = (long long)args.g.row_offsets[row + 1];
        }
        Value acc = ReduceIdentity<Value>(red.op);
        for (long long p0 = begin; p0 < end; p0 += 32) {           // uniform trip count
            const long long p = p0 + lane;
            bool ok = false;
            VertexId nbr = -1;
            if (p < end) {
                nbr = util::LoadStream(args.g.col_indices + p);
                ++my_edges;
                const VertexId s_id = Problem::MARK_PREDECESSORS ? v : (VertexId)args.iteration;
                ok = VisitEdge<Policy, Problem, Functor>(&ds, s_id, nbr, (SizeT)p, (SizeT)i);
                if (ok && red.rtype != EMPTY)
                    acc = ReduceCombine<Value>(red.op, acc, red.values[red.rtype == VERTEX ? (long long)nbr : p]);
            }
            if (args.output == OUT_COMPACT) {
                const unsigned m = __ballot_sync(util::kFullMask, ok);
                if (m) {
                    const int leader = __ffs(m) - 1;
                    unsigned long long b = 0;
                    if (lane == leader) b = atomicAdd(&slot_out->out_len, (unsigned long long)__popc(m));
                    b = __shfl_sync(util::kFullMask, b, leader);
                    if (ok) {
                        const unsigned long long at = b + __popc(m & ((1u << lane) - 1));
                        if (at < args.out_cap) args.out_queue[at] = OutputIsEdges(args.type) ? (VertexId)p : nbr;
                        else slot_out->overflow = 1;
                    }
                }
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) acc = ReduceCombine<Value>(red.op, acc, __shfl_xor_sync(util::kFullMask, acc, o));
        if (lane == 0 && red.reduced) red.reduced[i] = acc;
    }
    BlockAccumulate<THREADS>(my_edges, s_red, &slot_out->edges);
}

// reference signature: advance::LaunchKernel<..>(..., d_value_to_reduce, d_reduce_frontier, d_reduced_value, R_OP, R_TYPE)
template <typename Policy, typename Problem, typename Functor, typename Value>
cudaError_t LaunchNeighborReduce(int sm_count, const typename Problem::DataSlice &ds,
                                 const Args<typename Policy::VertexId, typename Policy::SizeT> &args,
                                 const ReduceArgs<Value> &red, cudaStream_t stream)
{
    if (args.output == OUT_TOMBSTONE)
        return util::GRError(cudaErrorInvalidValue, "neighbour-reduce: compacted or no output only", __FILE__, __LINE__);
    const unsigned long long resident = (unsigned long long)sm_count * Policy::MIN_CTAS;
    unsigned long long ctas = (args.in_len + Policy::THREADS / 32 - 1) / (Policy::THREADS / 32);
    const unsigned grid = (unsigned)(ctas < 1 ? 1 : (ctas < resident ? ctas : resident));
    NeighborReduceKernel<Policy, Problem, Functor, Value><<<grid, Policy::THREADS, 0, stream>>>(ds, args, red);
    return util::GRError(cudaGetLastError(), "advance NeighborReduceKernel launch failed", __FILE__, __LINE__);
}

}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
