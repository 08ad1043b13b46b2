// cc_capi.cu -- C entry points for connected components: the reference's gunrock_cc_func
// (gunrock/gunrock.h:119-124; implementation replaced: gunrock/app/cc/cc_app.cu:121-285)
// and the resident-graph grb_cc of include/gunrock_b200.h.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock/gunrock.h"
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/app/cc/cc_enactor.cuh"

// counterpart of run_cc<VertexId, Value, SizeT> (cc_app.cu:121-190): Init -> Reset -> Enact -> Extract
template <typename SizeT>
static cudaError_t RunCc(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t stream,
                         int *labels_host, int64_t *num_components, grb_stats *stats)
{
    typedef gunrock::app::cc::CCProblem<int, SizeT, int> Problem;
    typedef gunrock::app::cc::CCEnactor<Problem> Enactor;
    if (g->part.log_parts != 0)
        return gunrock::util::GRError(cudaErrorNotSupported, "CC runs on unpartitioned graphs only", __FILE__, __LINE__);
    Problem problem;
    Enactor enactor;
    cudaEvent_t e0, e1, e2;
    GR_TRY(cudaEventCreate(&e0)); GR_TRY(cudaEventCreate(&e1)); GR_TRY(cudaEventCreate(&e2));
    cudaError_t err = cudaSuccess;
    do {
        if ((err = problem.Init(g, stream)) != cudaSuccess) break;
        if ((err = enactor.Setup(&problem, stream)) != cudaSuccess) break;
        if ((err = cudaEventRecord(e0, stream)) != cudaSuccess) break;
        if ((err = problem.Reset()) != cudaSuccess) break;
        if ((err = cudaEventRecord(e1, stream)) != cudaSuccess) break;
        if ((err = enactor.Enact(&problem)) != cudaSuccess) break;
        if ((err = cudaEventRecord(e2, stream)) != cudaSuccess) break;
        if ((err = cudaEventSynchronize(e2)) != cudaSuccess) break;
        if ((err = problem.Extract(labels_host)) != cudaSuccess) break;
        if (num_components) *num_components = (int64_t)problem.num_components;
        if (stats) {
            float r = 0, t = 0;
            cudaEventElapsedTime(&r, e0, e1); cudaEventElapsedTime(&t, e1, e2);
            memset(stats, 0, sizeof(*stats));
            stats->reset_ms = r; stats->elapsed_ms = t;
            stats->iterations = enactor.enactor_stats.iteration;
            stats->edges_inspected = enactor.enactor_stats.edges_inspected;
            stats->edges_visited = (int64_t)enactor.enactor_stats.edges_inspected;
            stats->total_queued = enactor.enactor_stats.total_queued;
            stats->nodes_visited = g->local_nodes;
            stats->algo_bytes = enactor.enactor_stats.algo_bytes;
            stats->push_levels = 1;
            stats->gpu_launches = enactor.enactor_stats.launches;
        }
    } while (0);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    return err;
}

extern "C" int grb_cc(grb_graph_t g, int32_t *labels_host, int64_t *num_components, grb_stats *stats)
{
    if (!g) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    return (int)(g->g32 ? RunCc<int>(g->g32, g->stream, labels_host, num_components, stats)
                        : RunCc<long long>(g->g64, g->stream, labels_host, num_components, stats));
}

extern "C" void gunrock_cc_func(struct GunrockGraph *graph_out, const struct GunrockGraph *graph_in,
                                struct GunrockConfig configs, struct GunrockDataType data_type)
{
    // only <int, int, int> exists in the reference (cc_app.cu:216-262)
    if (data_type.VTXID_TYPE != VTXID_INT || data_type.SIZET_TYPE != SIZET_INT || data_type.VALUE_TYPE != VALUE_INT) {
        printf("Not Yet Support This DataType Combination.\n");
        return;
    }
    const long long n = (long long)graph_in->num_nodes, m = (long long)graph_in->num_edges;
    grb_graph_t g = nullptr;
    // test_cc.c:21-22 sets only configs.device; nothing else of `configs` is read
    if (grb_graph_from_host(configs.device, n, m, graph_in->row_offsets, 4, (const int32_t *)graph_in->col_indices,
                            nullptr, 0, 4, &g))
        return;
    int *h_labels = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));      // caller frees (cc_app.cu:143,182)
    int64_t comps = 0;
    grb_stats st;
    memset(&st, 0, sizeof(st));
    const int rc = grb_cc(g, h_labels, &comps, &st);
    if (rc == 0) {
        printf("[CC] finished. \n elapsed: %.4f ms\n num_components: %lld\n", st.elapsed_ms, (long long)comps);
        graph_out->node_values = h_labels;
    } else {
        free(h_labels);
        graph_out->node_values = nullptr;
    }
    grb_graph_destroy(g);
    cudaDeviceSynchronize();
}
