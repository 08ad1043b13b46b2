// kernel_policy.cuh -- compile-time tuning and argument blocks of the advance operator.
//
// Counterpart of the reference's oprtr/advance/kernel_policy.cuh:43-174 (MODE /
// TYPE enums + composite KernelPolicy tuned for sm_30) and of the parameter
// lists of advance::LaunchKernel (oprtr/advance/kernel.cuh:101-129).  The
// tuning here is for sm_100a: 148 SMs, 64 warps and 227 KB of shared memory
// per SM, 126 MB L2.
#pragma once
#include <cstdint>
#include "../../graph/csr.cuh"
#include "../work_progress.cuh"

namespace gunrock {
namespace oprtr {
namespace advance {

// load-balancing strategy (names follow the reference's advance::MODE)
enum MODE {
    TWC_FORWARD,    // thread / warp / CTA cooperative gather, one kernel, best for small frontiers
    LB,             // merge-path partitioned gather over the scanned frontier degrees
    TWC_BACKWARD,   // pull direction (bottom-up) over the inverse graph
    LB_BACKWARD     // alias of the pull direction kept for source compatibility
};

// what the input / output frontier holds (reference: advance::TYPE, oprtr/advance/kernel_policy.cuh:54-60):
// an input EDGE id stands for the vertex that edge points to (edge_map_partitioned/kernel.cuh:356-359,
// 385-388); an output edge frontier receives the CSR entry index of every successful edge instead of
// its destination (:409-413).  Edge ids travel in VertexId queues, as in the reference (SizeT = int).
enum TYPE { V2V, V2E, E2V, E2E };
__host__ __device__ __forceinline__ bool InputIsEdges(TYPE t)  { return t == E2V || t == E2E; }
__host__ __device__ __forceinline__ bool OutputIsEdges(TYPE t) { return t == V2E || t == E2E; }

// how successful edges are written
enum OUTPUT {
    OUT_NONE,       // functor side effects only (reference: d_out == NULL, pr_enactor.cuh:363)
    OUT_COMPACT,    // only destinations whose CondEdge passed, densely packed
    OUT_TOMBSTONE   // reference layout: one slot per inspected edge, -1 where CondEdge failed
};

template <
    typename _Problem,
    int _THREADS          = 256,   // CTA size of the gather kernels
    int _EDGES_PER_THREAD = 4,     // LB: neighbour entries per thread per tile
    int _MIN_CTAS         = 4,     // resident CTAs per SM the kernels are bounded for
    bool _PRE_CULL        = false, // operator probes Problem's visited bitmask before CondEdge
    bool _TRACK_OUT_DEGREE = false,// accumulate sum of out-degrees of emitted vertices
    int _WARP_THRESHOLD   = 32,    // TWC: lists >= this are gathered by a warp
    int _CTA_THRESHOLD    = 1024>  // TWC: lists >= this are gathered by the whole CTA
struct KernelPolicy {
    typedef _Problem Problem;
    typedef typename Problem::VertexId VertexId;
    typedef typename Problem::SizeT    SizeT;
    // type the scanned frontier degrees are held in
    typedef unsigned long long ScanT;

    static const int  THREADS          = _THREADS;
    static const int  WARPS            = _THREADS / 32;
    static const int  EDGES_PER_THREAD = _EDGES_PER_THREAD;
    static const int  TILE_EDGES       = _THREADS * _EDGES_PER_THREAD;
    static const int  WINDOW           = TILE_EDGES + 1;    // frontier items staged per tile
    static const int  STAGE_CAP        = 2 * TILE_EDGES;    // CTA-local enqueue staging buffer
    static const int  MIN_CTAS         = _MIN_CTAS;
    static const bool PRE_CULL         = _PRE_CULL;
    static const bool TRACK_OUT_DEGREE = _TRACK_OUT_DEGREE;
    static const int  WARP_THRESHOLD   = _WARP_THRESHOLD;
    static const int  CTA_THRESHOLD    = _CTA_THRESHOLD;
    // degree scan: items per CTA of the single-pass look-back scan
    static const int  SCAN_THREADS     = 512;
    static const int  SCAN_ROUNDS      = 2;
    static const int  SCAN_TILE        = SCAN_THREADS * SCAN_ROUNDS;
};

// per-call arguments common to all advance modes (passed in the parameter bank)
template <typename VertexId, typename SizeT>
struct Args {
    graph::CsrView<VertexId, SizeT> g;          // graph to walk (inverse view for pull)
    const VertexId     *in_queue;               // input frontier (vertex ids, -1 = hole); NULL = every owned
                                                // vertex in local-row order (in_len = local_nodes)
    VertexId           *out_queue;              // output frontier or NULL
    unsigned long long  in_len;                 // host-known input length ...
    const unsigned long long *d_in_len;         // ... or, if non-NULL, read it here
    unsigned long long  out_cap;                // capacity of out_queue
    LevelCounters      *slots;                  // the 4 rotating counter slots
    long long           queue_index;            // accumulate into slot queue_index+1
    long long           iteration;              // passed as s_id when !MARK_PREDECESSORS
    OUTPUT              output;
    TYPE                type = V2V;
};

// scratch the LB mode needs; sized once for the largest frontier (no per-call cudaMalloc,
// unlike d_scanned_edges in the reference: app/bfs/bfs_enactor.cuh:254-258)
template <typename SizeT>
struct LbScratch {
    util::DeviceArray<unsigned long long> scanned;     // [max_items + 1] exclusive degree scan
    util::DeviceArray<SizeT>              row_start;   // [max_items]     first CSR entry of the item
    util::DeviceArray<unsigned long long> tile_desc;   // look-back descriptors of the scan
    util::DeviceArray<long long>          part_start;  // [max_tiles + 1] first item of each edge tile
    util::DeviceArray<unsigned>           ticket;      // scan tiles are handed out in arrival order
    util::DeviceArray<unsigned>           mask_before; // streaming instance: the visited mask before the level (lazily sized)
    unsigned epoch = 0;                                // tags descriptors => no memset per call

    cudaError_t Allocate(size_t max_items, size_t max_edges, int scan_tile, int tile_edges, cudaStream_t s)
    {
        GR_TRY(scanned.Allocate(max_items + 1));
        GR_TRY(row_start.Allocate(max_items + 1));
        GR_TRY(tile_desc.Allocate(max_items / scan_tile + 2));
        GR_TRY(part_start.Allocate(max_edges / tile_edges + 2));
        GR_TRY(ticket.Allocate(1));
        GR_TRY(ticket.Zero(s));
        epoch = 0;
        return tile_desc.Zero(s);
    }
};

template <typename SizeT>
struct LbScratchView {
    unsigned long long *scanned;
    SizeT              *row_start;
    unsigned long long *tile_desc;
    long long          *part_start;
    unsigned           *ticket;
    unsigned long long  part_cap;       // entries in part_start
    unsigned            epoch;
};

}  // namespace advance
}  // namespace oprtr
}  // namespace gunrock
