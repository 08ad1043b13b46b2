// sssp_capi.cu -- C entry points for SSSP: the reference's gunrock_sssp_func
// (gunrock/gunrock.h:126-131; implementation replaced: gunrock/app/sssp/sssp_app.cu:139-397)
// and the resident-graph grb_sssp of include/gunrock_b200.h.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock/gunrock.h"
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "source_select.h"
#include "../gunrock/app/sssp/sssp_enactor.cuh"

// counterpart of run_sssp<VertexId, Value, SizeT, MARK_PREDECESSORS> (sssp_app.cu:139-227):
// Init -> Reset -> Enact -> Extract
template <typename SizeT, bool MARK_PATHS>
static cudaError_t RunSssp(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t stream, int src, int delta_factor,
                           unsigned *dist_host, int *preds_host, grb_stats *stats)
{
    typedef gunrock::app::sssp::SSSPProblem<int, SizeT, MARK_PATHS> Problem;
    typedef gunrock::app::sssp::SSSPEnactor<Problem> Enactor;
    Problem problem;
    Enactor enactor;
    cudaEvent_t e0, e1, e2;
    GR_TRY(cudaEventCreate(&e0)); GR_TRY(cudaEventCreate(&e1)); GR_TRY(cudaEventCreate(&e2));
    cudaError_t err = cudaSuccess;
    do {
        if ((err = problem.Init(g, delta_factor, stream)) != cudaSuccess) break;
        if ((err = cudaEventRecord(e0, stream)) != cudaSuccess) break;
        if ((err = problem.Reset(src, 1.0)) != cudaSuccess) break;
        if ((err = enactor.Setup(&problem, stream)) != cudaSuccess) break;
        if ((err = cudaEventRecord(e1, stream)) != cudaSuccess) break;
        if ((err = enactor.Enact(&problem, src)) != cudaSuccess) break;
        if ((err = cudaEventRecord(e2, stream)) != cudaSuccess) break;
        if ((err = cudaEventSynchronize(e2)) != cudaSuccess) break;
        if (dist_host || preds_host) err = problem.Extract(dist_host, MARK_PATHS ? preds_host : nullptr);
        if (stats) {
            float r = 0, t = 0;
            cudaEventElapsedTime(&r, e0, e1); cudaEventElapsedTime(&t, e1, e2);
            memset(stats, 0, sizeof(*stats));
            stats->reset_ms = r; stats->elapsed_ms = t;
            stats->iterations = enactor.enactor_stats.iteration;
            stats->edges_inspected = enactor.relaxations;
            stats->edges_visited = (int64_t)enactor.relaxations;
            stats->total_queued = enactor.enactor_stats.total_queued;
            stats->nodes_visited = (int64_t)enactor.enactor_stats.total_queued;
            stats->algo_bytes = enactor.enactor_stats.algo_bytes;
            stats->push_levels = (int32_t)enactor.enactor_stats.iteration;
            stats->pull_levels = enactor.bucket_jumps;
            stats->gpu_launches = enactor.enactor_stats.launches;
            stats->reserved = (int32_t)problem.delta;
        }
    } while (0);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    return err;
}

extern "C" int grb_sssp(grb_graph_t g, int32_t src, int32_t delta_factor, int32_t mark_pred,
                        uint32_t *dist_host, int32_t *preds_host, grb_stats *stats)
{
    if (!g) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    cudaError_t e;
    if (g->g32) e = mark_pred ? RunSssp<int, true>(g->g32, g->stream, src, delta_factor, dist_host, preds_host, stats)
                              : RunSssp<int, false>(g->g32, g->stream, src, delta_factor, dist_host, preds_host, stats);
    else        e = mark_pred ? RunSssp<long long, true>(g->g64, g->stream, src, delta_factor, dist_host, preds_host, stats)
                              : RunSssp<long long, false>(g->g64, g->stream, src, delta_factor, dist_host, preds_host, stats);
    return (int)e;
}

extern "C" void gunrock_sssp_func(struct GunrockGraph *graph_out, void *predecessor,
                                  const struct GunrockGraph *graph_in, struct GunrockConfig configs,
                                  struct GunrockDataType data_type)
{
    // only <int, unsigned int, int> exists in the reference (sssp_app.cu:250-345)
    if (data_type.VTXID_TYPE != VTXID_INT || data_type.SIZET_TYPE != SIZET_INT || data_type.VALUE_TYPE != VALUE_UINT) {
        printf("Not Yet Support This DataType Combination.\n");
        return;
    }
    const long long n = (long long)graph_in->num_nodes, m = (long long)graph_in->num_edges;
    const int *h_off = (const int *)graph_in->row_offsets;
    const int src = grb_pick_source(configs.src_mode, configs.src_node, n, h_off);
    grb_graph_t g = nullptr;
    if (grb_graph_from_host(configs.device, n, m, h_off, 4, (const int32_t *)graph_in->col_indices,
                            (const uint32_t *)graph_in->edge_values, 0, 4, &g))
        return;
    unsigned *h_labels = (unsigned *)malloc(sizeof(unsigned) * (size_t)(n > 0 ? n : 1));   // caller frees (:161,208)
    grb_stats st;
    memset(&st, 0, sizeof(st));
    const int rc = grb_sssp(g, src, configs.delta_factor, configs.mark_pred ? 1 : 0, h_labels,
                            configs.mark_pred ? (int32_t *)predecessor : nullptr, &st);
    if (rc == 0) {
        printf("[SSSP] finished. \n elapsed: %.4f ms\n src: %d\n iterations: %lld\n relaxations: %llu\n delta: %d\n",
               st.elapsed_ms, src, (long long)st.iterations, (unsigned long long)st.edges_inspected, st.reserved);
        graph_out->node_values = h_labels;
    } else {
        free(h_labels);
        graph_out->node_values = nullptr;
    }
    grb_graph_destroy(g);
    cudaDeviceSynchronize();
}
