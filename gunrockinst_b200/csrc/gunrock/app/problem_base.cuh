// problem_base.cuh -- what every primitive's Problem shares: the device graph
// slice and the ping-pong frontier queues.
//
// Counterpart of the reference's ProblemBase (gunrock/app/problem_base.cuh:
// 73-431).  Differences that matter on a 180 GB part: the graph slice can be
// BORROWED from a resident graph handle (no re-upload per run), and the two
// frontier queues are sized for what the fused operators can actually emit
// (n vertex ids, duplicates are impossible once the visited bitmask claims a
// vertex) instead of max(m, n) * queue_sizing ids each (:378-382) -- 0.27 GB
// instead of 8 GB at scale 25.  Problems that emit duplicates (idempotent BFS)
// ask for EDGE_FRONTIERS sizing explicitly.
#pragma once
#include "../graph/csr.cuh"
#include "../util/memset_kernel.cuh"

namespace gunrock {
namespace app {

enum FrontierType {
    VERTEX_FRONTIERS,   // O(n) ping-pong vertex frontiers
    EDGE_FRONTIERS,     // O(m) ping-pong frontiers (outputs may hold duplicates / tombstones)
    MIXED_FRONTIERS
};

template <typename _VertexId, typename _SizeT>
struct ProblemBase {
    typedef _VertexId VertexId;
    typedef _SizeT    SizeT;
    typedef graph::DeviceCsr<VertexId, SizeT> Graph;

    Graph       *graph = nullptr;     // borrowed or owned, see owns_graph
    bool         owns_graph = false;
    cudaStream_t stream = 0;

    // ping-pong frontier queues (selector picks the input)
    util::DeviceArray<VertexId> frontier_queues[2];
    unsigned long long          frontier_elements[2] = {0, 0};

    virtual ~ProblemBase() { if (owns_graph && graph) delete graph; }

    // run on a graph that already lives on the device
    cudaError_t InitBase(Graph *resident, cudaStream_t s)
    {
        if (owns_graph && graph) delete graph;
        graph = resident; owns_graph = false; stream = s;
        return cudaSuccess;
    }

    // upload a host CSR (the reference's Init path: problem_base.cuh:279-303)
    template <typename HostSizeT>
    cudaError_t InitBase(long long n, long long m, const HostSizeT *h_offsets, const VertexId *h_cols,
                         const unsigned *h_weights, int device, cudaStream_t s)
    {
        if (owns_graph && graph) delete graph;
        graph = new Graph(); owns_graph = true; stream = s;
        graph->device = device;
        return graph->FromHost(n, m, h_offsets, h_cols, h_weights, s);
    }

    // size the two frontier queues (reference: ProblemBase::Reset, :352-431)
    cudaError_t ResetBase(FrontierType type, double queue_sizing)
    {
        if (queue_sizing <= 0) queue_sizing = 1.0;
        unsigned long long n = (unsigned long long)graph->local_nodes, m = (unsigned long long)graph->edges;
        unsigned long long want = (type == VERTEX_FRONTIERS) ? n : (m > n ? m : n);
        want = (unsigned long long)((double)want * queue_sizing) + 1;
        for (int i = 0; i < 2; ++i) {
            GR_TRY(frontier_queues[i].Allocate((size_t)want));
            frontier_elements[i] = want;
        }
        return cudaSuccess;
    }
};

}  // namespace app
}  // namespace gunrock
