// pr_capi.cu -- C entry points for PageRank: the reference's gunrock_pr_func
// (gunrock/gunrock.h:133-140; implementation replaced: gunrock/app/pr/pr_app.cu:120-339)
// and the resident-graph grb_pagerank of include/gunrock_b200.h.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock/gunrock.h"
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "source_select.h"
#include "../gunrock/app/pr/pr_enactor.cuh"
#include "../gunrock/app/pr/mgpr_enactor.cuh"

// counterpart of run_page_rank<VertexId, Value, SizeT> (pr_app.cu:120-195)
template <typename SizeT>
static cudaError_t RunPr(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t stream, int src, double delta,
                         double threshold, int max_iter, double *ranks64_host, int *sorted_ids_host,
                         float *sorted_ranks_host, size_t top, grb_stats *stats)
{
    typedef gunrock::app::pr::PRProblem<int, SizeT, double> Problem;
    typedef gunrock::app::pr::PREnactor<Problem> Enactor;
    if (g->part.log_parts != 0)
        return gunrock::util::GRError(cudaErrorNotSupported, "PageRank runs on unpartitioned graphs only", __FILE__, __LINE__);
    Problem problem;
    Enactor enactor;
    cudaEvent_t e0, e1, e2;
    GR_TRY(cudaEventCreate(&e0)); GR_TRY(cudaEventCreate(&e1)); GR_TRY(cudaEventCreate(&e2));
    cudaError_t err = cudaSuccess;
    do {
        if ((err = problem.Init(g, stream)) != cudaSuccess) break;
        if ((err = enactor.Setup(&problem, stream)) != cudaSuccess) break;
        if ((err = cudaEventRecord(e0, stream)) != cudaSuccess) break;
        if ((err = problem.Reset(src, delta, threshold)) != cudaSuccess) break;
        if ((err = cudaEventRecord(e1, stream)) != cudaSuccess) break;
        if ((err = enactor.Enact(&problem, max_iter)) != cudaSuccess) break;
        if ((err = cudaEventRecord(e2, stream)) != cudaSuccess) break;
        if ((err = cudaEventSynchronize(e2)) != cudaSuccess) break;
        if ((err = problem.Extract(ranks64_host, sorted_ids_host, sorted_ranks_host, top)) != cudaSuccess) break;
        if (stats) {
            float r = 0, t = 0;
            cudaEventElapsedTime(&r, e0, e1); cudaEventElapsedTime(&t, e1, e2);
            memset(stats, 0, sizeof(*stats));
            stats->reset_ms = r; stats->elapsed_ms = t;
            stats->iterations = enactor.enactor_stats.iteration;
            stats->edges_inspected = enactor.enactor_stats.edges_inspected;
            stats->edges_visited = (int64_t)enactor.enactor_stats.edges_inspected;
            stats->total_queued = enactor.enactor_stats.total_queued;
            stats->nodes_visited = (int64_t)enactor.valid_nodes;
            stats->algo_bytes = enactor.enactor_stats.algo_bytes;
            stats->push_levels = (int32_t)enactor.enactor_stats.iteration;
            stats->pull_levels = enactor.prune_sweeps;
            stats->gpu_launches = enactor.enactor_stats.launches;
            stats->reserved = (int32_t)(enactor.sweep_ms * 1000.0f);      // microseconds of the power iteration alone
        }
    } while (0);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    return err;
}

// out[v] = in[to_new[v]]: ranks computed on the degree-ordered twin, back in the caller's vertex order
static __global__ void RanksThroughKernel(const double *in, const int *to_new, long long nodes, double *out)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nodes; v += stride) out[v] = in[to_new[v]];
}

// Symmetric graphs: the gather formulation of mgpr_enactor.cuh on one device (in-edges == the rows
// already held, so a sweep sums contributions from a dense vector instead of issuing one fp64 atomic
// per edge to a random line: 11.4 instead of 16.8 ms per sweep at scale 25).  Same pruning, damping,
// moving test and sweep count; outputs as RunPr.
// to_new != nullptr: g is the degree-ordered twin of the caller's graph (graph/build.cuh) and to_new[v] the
// twin's id of the caller's vertex v -- the same graph with the hubs' contributions packed into the first
// few MB of the vector every sweep gathers from; src goes in and the ranks come out in the caller's ids.
template <typename SizeT>
static cudaError_t RunPrGather(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t stream, int src, double delta,
                               double threshold, int max_iter, double *ranks64_host, int *sorted_ids_host,
                               float *sorted_ranks_host, size_t top, grb_stats *stats, const int *to_new = nullptr)
{
    typedef gunrock::app::pr::MGPREnactor<SizeT> Enactor;
    Enactor enactor;
    typename Enactor::RunStats st;
    const size_t n = (size_t)g->local_nodes;
    if (to_new && src >= 0 && (size_t)src < n) {
        GR_TRY(cudaMemcpyAsync(&src, to_new + src, sizeof(int), cudaMemcpyDeviceToHost, stream));
        GR_TRY(cudaStreamSynchronize(stream));
    }
    GR_TRY(enactor.Setup(g, stream));
    GR_TRY(enactor.Run(src, delta, threshold, max_iter, &st));
    gunrock::util::DeviceArray<double> caller_order;
    const double *d_rank = enactor.rank.ptr;
    if (to_new) {
        GR_TRY(caller_order.Allocate(n + 1));
        RanksThroughKernel<<<gunrock::util::FillGrid(n), 256, 0, stream>>>(enactor.rank.ptr, to_new, (long long)n, caller_order.ptr);
        GR_TRY(cudaGetLastError());
        d_rank = caller_order.ptr;
    }
    if (ranks64_host) {
        if (n) GR_TRY(cudaMemcpyAsync(ranks64_host, d_rank, sizeof(double) * n, cudaMemcpyDeviceToHost, stream));
        GR_TRY(cudaStreamSynchronize(stream));
    }
    if (sorted_ids_host || sorted_ranks_host) {
        // all (id, fp32 rank) pairs in descending rank, ties by ascending id (pr_enactor.cuh:497-500)
        gunrock::util::DeviceArray<unsigned long long> keys_a, keys_b;
        gunrock::util::DeviceArray<int> d_ids;
        gunrock::util::DeviceArray<float> d_ranks;
        gunrock::util::RadixSortScratch scratch;
        GR_TRY(keys_a.Allocate(n)); GR_TRY(keys_b.Allocate(n)); GR_TRY(d_ids.Allocate(n)); GR_TRY(d_ranks.Allocate(n));
        gunrock::app::pr::MakeSortKeysKernel<<<gunrock::util::FillGrid(n), 256, 0, stream>>>(d_rank, n, keys_a.ptr);
        GR_TRY(cudaGetLastError());
        unsigned long long *sorted = nullptr;
        GR_TRY(gunrock::util::RadixSort64(keys_a.ptr, keys_b.ptr, n, 64, scratch, &sorted, stream));
        gunrock::app::pr::UnpackSortedKernel<<<gunrock::util::FillGrid(n), 256, 0, stream>>>(sorted, n, d_ids.ptr, d_ranks.ptr);
        GR_TRY(cudaGetLastError());
        const size_t k = (top && top < n) ? top : n;
        if (sorted_ids_host)   GR_TRY(cudaMemcpyAsync(sorted_ids_host, d_ids.ptr, sizeof(int) * k, cudaMemcpyDeviceToHost, stream));
        if (sorted_ranks_host) GR_TRY(cudaMemcpyAsync(sorted_ranks_host, d_ranks.ptr, sizeof(float) * k, cudaMemcpyDeviceToHost, stream));
        GR_TRY(cudaStreamSynchronize(stream));
    }
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->elapsed_ms = (float)st.elapsed_ms;
        stats->iterations = st.iterations;
        stats->edges_inspected = st.edges;
        stats->edges_visited = (int64_t)st.edges;
        stats->total_queued = st.valid_nodes * (unsigned long long)st.iterations;
        stats->nodes_visited = (int64_t)st.valid_nodes;
        // per walked entry 4 (column) + 8 (contribution); per alive vertex and sweep: queue 4, offsets,
        // accumulator 8 + 8 (not with the row-binned sweep, whose sums stay in registers), rank 8 + 8,
        // degree 4, contribution 8
        stats->algo_bytes = 12ull * st.edges + (unsigned long long)st.iterations * st.valid_nodes *
                                                   ((st.by_rows ? 32ull : 48ull) + 2ull * sizeof(SizeT));
        stats->push_levels = (int32_t)st.iterations;
        stats->pull_levels = (int32_t)st.prune_sweeps;
        stats->gpu_launches = st.launches;
        stats->reserved = (int32_t)(st.sweep_ms * 1000.0);         // microseconds of the power iteration alone
    }
    GR_TRY(cudaStreamSynchronize(stream));      // caller_order is released below
    return cudaSuccess;
}

template <typename SizeT>
static cudaError_t RunPrAny(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t stream, int src, double delta,
                            double threshold, int max_iter, double *ranks64_host, int *sorted_ids_host,
                            float *sorted_ranks_host, size_t top, grb_stats *stats)
{
    if (g->symmetric && g->part.log_parts == 0 && g->local_nodes > 0 && g->edges > 0)
        return RunPrGather<SizeT>(g, stream, src, delta, threshold, max_iter, ranks64_host, sorted_ids_host,
                                  sorted_ranks_host, top, stats);
    return RunPr<SizeT>(g, stream, src, delta, threshold, max_iter, ranks64_host, sorted_ids_host,
                        sorted_ranks_host, top, stats);
}

static int PrDispatch(grb_graph_t g, int32_t src, double delta, double threshold, int32_t max_iter,
                      double *ranks64_host, int32_t *sorted_ids_host, float *sorted_ranks_host, size_t top,
                      grb_stats *stats)
{
    if (!g) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    // a symmetric graph that already has its degree-ordered twin attached (grb_graph_build_twin, or a search with
    // grb_bfs_config.degree_order) is swept on the twin -- the same multiset of entries under other vertex names
    // (BuildDegreeOrdered keeps duplicates and self loops; the entry counts are compared all the same);
    // GRB_PR_TWIN=0 keeps the caller's order
    const char *tw = getenv("GRB_PR_TWIN");
    if (g->has_twin() && !(tw && tw[0] == '0')) {
        if (g->twin32 && g->g32->symmetric && g->twin32->edges == g->g32->edges && g->g32->edges > 0)
            return (int)RunPrGather<int>(g->twin32, g->stream, src, delta, threshold, max_iter, ranks64_host,
                                         sorted_ids_host, sorted_ranks_host, top, stats, g->to_new.ptr);
        if (g->twin64 && g->g64->symmetric && g->twin64->edges == g->g64->edges && g->g64->edges > 0)
            return (int)RunPrGather<long long>(g->twin64, g->stream, src, delta, threshold, max_iter, ranks64_host,
                                               sorted_ids_host, sorted_ranks_host, top, stats, g->to_new.ptr);
    }
    return (int)(g->g32 ? RunPrAny<int>(g->g32, g->stream, src, delta, threshold, max_iter, ranks64_host,
                                        sorted_ids_host, sorted_ranks_host, top, stats)
                        : RunPrAny<long long>(g->g64, g->stream, src, delta, threshold, max_iter, ranks64_host,
                                              sorted_ids_host, sorted_ranks_host, top, stats));
}

extern "C" int grb_pagerank(grb_graph_t g, int32_t src, double delta, double threshold, int32_t max_iter,
                            double *ranks64_host, int32_t *sorted_ids_host, float *sorted_ranks_host,
                            grb_stats *stats)
{
    return PrDispatch(g, src, delta, threshold, max_iter, ranks64_host, sorted_ids_host, sorted_ranks_host, 0, stats);
}

extern "C" void gunrock_pr_func(struct GunrockGraph *graph_out, void *node_ids, void *page_rank,
                                const struct GunrockGraph *graph_in, struct GunrockConfig configs,
                                struct GunrockDataType data_type)
{
    // only <int, float, int> exists in the reference (pr_app.cu:207-300)
    if (data_type.VTXID_TYPE != VTXID_INT || data_type.SIZET_TYPE != SIZET_INT || data_type.VALUE_TYPE != VALUE_FLOAT) {
        printf("Not Yet Support This DataType Combination.\n");
        return;
    }
    (void)graph_out;        // the reference leaves graph_out untouched for PageRank
    const long long n = (long long)graph_in->num_nodes, m = (long long)graph_in->num_edges;
    const int *h_off = (const int *)graph_in->row_offsets;
    // pr_app.cu:232-258: any other mode value means "global" PageRank (src = -1)
    const int src = grb_pick_source(configs.src_mode, configs.src_node, n, h_off, -1);
    grb_graph_t g = nullptr;
    if (grb_graph_from_host(configs.device, n, m, h_off, 4, (const int32_t *)graph_in->col_indices, nullptr, 0, 4, &g))
        return;
    // The reference writes all n pairs (pr_problem.cuh:162-176) although test_pr.c:47-48 sizes the
    // buffers by top_nodes; writing min(n, top_nodes) is output compatible and cannot overrun.
    const size_t top = configs.top_nodes > 0 ? (size_t)configs.top_nodes : 0;
    grb_stats st;
    memset(&st, 0, sizeof(st));
    const int rc = PrDispatch(g, src, (double)configs.delta, (double)configs.error, configs.max_iter, nullptr,
                              (int32_t *)node_ids, (float *)page_rank, top, &st);
    if (rc == 0)
        printf("[PR] finished. \n elapsed: %.4f ms\n src: %d\n iterations: %lld\n", st.elapsed_ms, src,
               (long long)st.iterations);
    grb_graph_destroy(g);
    cudaDeviceSynchronize();
}
