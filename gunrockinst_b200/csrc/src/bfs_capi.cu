// bfs_capi.cu -- C entry points for BFS: the reference's gunrock_bfs_func and the
// resident-graph variants of include/gunrock_b200.h.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock/gunrock.h"
#include "bfs_runner.cuh"
#include "source_select.h"
#include <chrono>

static double grb_wall_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct grb_bfs_s {
    grb_graph_t graph = nullptr;
    BfsRunnerBase *runner = nullptr;
    ~grb_bfs_s() { delete runner; }
};

extern "C" int grb_bfs_create(grb_graph_t g, const grb_bfs_config *cfg, grb_bfs_t *out)
{
    if (!g || !out) return (int)cudaErrorInvalidValue;
    *out = nullptr;
    GRB_TRY(cudaSetDevice(g->device()));
    cudaError_t err = cudaSuccess;
    const bool single = (g->g32 ? g->g32->part.log_parts : g->g64->part.log_parts) == 0;
    const bool mapped = g->g32 ? g->g32->streams_from_host : g->g64->streams_from_host;
    const bool symmetric = g->g32 ? g->g32->symmetric : g->g64->symmetric;
    if (cfg && cfg->direction_optimized && !symmetric) {
        // the pull direction walks in-edges: a directed graph needs its transpose (the reference's DOBFS
        // takes an inverse graph, app/dobfs/dobfs_problem.cuh:211-260); built once per graph handle
        if (single && !mapped) GRB_TRY(g->EnsureTranspose());
        else fprintf(stderr, "libgunrock (B200 build): direction-optimised BFS on a directed graph that is %s: "
                             "no inverse graph can be built, the search runs push-only\n",
                     single ? "read in place from host memory" : "partitioned");
    }
    const bool twin = cfg && cfg->degree_order && single && !mapped;
    if (twin) GRB_TRY(g->EnsureTwin());
    grb_bfs_s *c = new grb_bfs_s();
    c->graph = g;
    if (twin)
        c->runner = g->g32 ? MakeBfsRunner32(g->twin32, cfg, g->stream, &err, g->to_new.ptr, g->to_old.ptr)
                           : MakeBfsRunner64(g->twin64, cfg, g->stream, &err, g->to_new.ptr, g->to_old.ptr);
    else
        c->runner = g->g32 ? MakeBfsRunner32(g->g32, cfg, g->stream, &err) : MakeBfsRunner64(g->g64, cfg, g->stream, &err);
    if (!c->runner) { delete c; return (int)(err ? err : cudaErrorUnknown); }
    *out = c;
    return 0;
}

extern "C" int grb_bfs_run(grb_bfs_t ctx, int32_t src, int32_t *labels_host, int32_t *preds_host, grb_stats *stats)
{
    if (!ctx) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(ctx->graph->device()));
    GRB_TRY(ctx->runner->Run(src, labels_host, preds_host, stats));
    return 0;
}

extern "C" int grb_bfs_run_many(grb_bfs_t ctx, const int32_t *srcs, int32_t count, grb_stats *stats)
{
    if (!ctx || (!srcs && count > 0)) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(ctx->graph->device()));
    GRB_TRY(ctx->runner->RunMany(srcs, count, stats));
    return 0;
}

extern "C" int grb_bfs_device_labels(grb_bfs_t ctx, void **labels_dev)
{
    if (!ctx || !labels_dev) return (int)cudaErrorInvalidValue;
    *labels_dev = ctx->runner->DeviceLabels();
    return 0;
}

extern "C" int grb_bfs_levels(grb_bfs_t ctx, grb_level_record *out, int32_t cap, int32_t *count)
{
    if (!ctx || !count) return (int)cudaErrorInvalidValue;
    const auto &lv = ctx->runner->Levels();
    *count = (int32_t)lv.size();
    for (int i = 0; i < (int)lv.size() && i < cap && out; ++i) {
        out[i].level = lv[i].level; out[i].kind = lv[i].pull; out[i].frontier = lv[i].frontier;
        out[i].edges = lv[i].edges; out[i].discovered = lv[i].discovered; out[i].candidates = lv[i].candidates;
        out[i].ms = lv[i].ms; out[i].switch_ms = lv[i].switch_ms;
    }
    return 0;
}

extern "C" void grb_bfs_destroy(grb_bfs_t ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->graph->device());
    delete ctx;
}

// ---------------------------------------------------------------------------------------
// reference entry point (gunrock/gunrock.h:106-110; implementation it replaces:
// gunrock/app/bfs/bfs_app.cu:241-402)
// ---------------------------------------------------------------------------------------
extern "C" void gunrock_bfs_func(struct GunrockGraph *graph_out, const struct GunrockGraph *graph_in,
                                 struct GunrockConfig configs, struct GunrockDataType data_type)
{
    if (data_type.VTXID_TYPE != VTXID_INT || data_type.SIZET_TYPE != SIZET_INT || data_type.VALUE_TYPE != VALUE_INT) {
        printf("Not Yet Support This DataType Combination.\n");       // bfs_app.cu:354-365
        return;
    }
    const long long n = (long long)graph_in->num_nodes, m = (long long)graph_in->num_edges;
    const int *h_off = (const int *)graph_in->row_offsets;
    const int src = grb_pick_source(configs.src_mode, configs.src_node, n, h_off);

    // GRB_TRACE=1: wall-clock split of the call on stderr (upload / context / search+download / teardown)
    const bool trace = getenv("GRB_TRACE") != nullptr;
    const double t_begin = trace ? grb_wall_ms() : 0.0;
    grb_graph_t g = nullptr;
    if (grb_graph_from_host(configs.device, n, m, h_off, 4, (const int32_t *)graph_in->col_indices, nullptr, 0, 4, &g))
        return;
    const double t_upload = trace ? grb_wall_ms() : 0.0;
    grb_bfs_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.mark_pred = configs.mark_pred; cfg.idempotence = configs.idempotence;
    cfg.queue_sizing = configs.queue_size;
    grb_bfs_t ctx = nullptr;
    if (grb_bfs_create(g, &cfg, &ctx)) { grb_graph_destroy(g); return; }

    const double t_ctx = trace ? grb_wall_ms() : 0.0;
    int *h_labels = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));     // caller frees (bfs_app.cu:165,211)
    int *h_preds = configs.mark_pred ? (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1)) : nullptr;
    grb_stats st;
    memset(&st, 0, sizeof(st));
    const int rc = grb_bfs_run(ctx, src, h_labels, h_preds, &st);
    if (rc == 0) {
        // same figures the reference prints (bfs_app.cu:103-118)
        const double rate = st.elapsed_ms > 0 ? (double)st.edges_visited / (st.elapsed_ms * 1000.0) : 0.0;
        printf("[BFS] finished. \n elapsed: %.4f ms\n src: %lld\n nodes_visited: %lld\n edges_visited: %lld\n"
               " rate: %.4f MiEdges/s\n search_depth: %lld\n",
               st.elapsed_ms, (long long)src, (long long)st.nodes_visited, (long long)st.edges_visited,
               rate, (long long)st.iterations);
        graph_out->node_values = h_labels;
    } else {
        free(h_labels);
        graph_out->node_values = nullptr;
    }
    const double t_run = trace ? grb_wall_ms() : 0.0;
    free(h_preds);                          // the C API never returns BFS predecessors (bfs_app.cu:166-170,227)
    grb_bfs_destroy(ctx);
    grb_graph_destroy(g);
    cudaDeviceSynchronize();                // bfs_app.cu:229
    if (trace)
        fprintf(stderr, "[grb trace] gunrock_bfs_func: upload %.3f ms, context %.3f ms, search+download %.3f ms "
                        "(device %.3f ms), teardown %.3f ms\n", t_upload - t_begin, t_ctx - t_upload, t_run - t_ctx,
                (double)(st.elapsed_ms + st.reset_ms), grb_wall_ms() - t_run);
}
