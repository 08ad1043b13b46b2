// bc_capi.cu -- C entry points for betweenness centrality and top-k degree centrality: the
// reference's gunrock_bc_func (gunrock/gunrock.h:112-116; gunrock/app/bc/bc_app.cu:46-271) and
// gunrock_topk_func (gunrock.h:133-141; gunrock/app/topk/topk_app.cu:157-352), plus the
// resident-graph grb_bc / grb_topk of include/gunrock_b200.h.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock/gunrock.h"
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "source_select.h"
#include "../gunrock/app/bc/bc_enactor.cuh"
#include "../gunrock/app/topk/topk_enactor.cuh"

template <typename T>
static __global__ void ScaleKernel(T *v, T f, long long n)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) v[i] *= f;
}

// counterpart of run_bc (bc_app.cu:46-139): src == -1 accumulates over every source, then the
// values are halved (MemsetScaleKernel(..., 0.5f), :111-112)
template <typename SizeT>
static cudaError_t RunBc(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t stream, int src,
                         float *bc_host, float *sigmas_host, int *labels_host, grb_stats *stats)
{
    typedef gunrock::app::bc::BCProblem<int, SizeT, float> Problem;
    typedef gunrock::app::bc::BCEnactor<Problem> Enactor;
    Problem problem;
    Enactor enactor;
    cudaEvent_t e0, e1;
    GR_TRY(cudaEventCreate(&e0)); GR_TRY(cudaEventCreate(&e1));
    cudaError_t err = cudaSuccess;
    grb_stats st;
    memset(&st, 0, sizeof(st));
    do {
        if ((err = problem.Init(g, stream)) != cudaSuccess) break;
        if ((err = enactor.Setup(&problem, stream)) != cudaSuccess) break;
        const long long n = g->nodes;
        const long long first = src < 0 ? 0 : src, last = src < 0 ? n : (long long)src + 1;
        if ((err = cudaEventRecord(e0, stream)) != cudaSuccess) break;
        for (long long s = first; s < last && err == cudaSuccess; ++s) {
            if ((err = problem.Reset((int)s)) != cudaSuccess) break;
            if ((err = enactor.Enact(&problem, (int)s)) != cudaSuccess) break;
            st.iterations += enactor.enactor_stats.iteration;
            st.edges_inspected += enactor.enactor_stats.edges_inspected;
            st.total_queued += enactor.enactor_stats.total_queued;
            st.algo_bytes += enactor.enactor_stats.algo_bytes;
            st.gpu_launches += enactor.enactor_stats.launches;
        }
        if (err != cudaSuccess) break;
        if (n > 0) {
            ScaleKernel<float><<<gunrock::util::FillGrid((unsigned long long)n), 256, 0, stream>>>(
                problem.bc_values.ptr, 0.5f, n);
            if ((err = cudaGetLastError()) != cudaSuccess) break;
            ++st.gpu_launches;
        }
        if ((err = cudaEventRecord(e1, stream)) != cudaSuccess) break;
        if ((err = cudaEventSynchronize(e1)) != cudaSuccess) break;
        err = problem.Extract(sigmas_host, bc_host, labels_host);
        float t = 0;
        cudaEventElapsedTime(&t, e0, e1);
        st.elapsed_ms = t;
        st.edges_visited = (int64_t)st.edges_inspected;
        st.nodes_visited = (int64_t)st.total_queued;
        if (stats) *stats = st;
    } while (0);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return err;
}

extern "C" int grb_bc(grb_graph_t g, int32_t src, float *bc_host, float *sigmas_host, int32_t *labels_host,
                      grb_stats *stats)
{
    if (!g) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    if (g->nodes() != g->local_nodes()) return (int)cudaErrorNotSupported;      // single-device graphs only
    return (int)(g->g32 ? RunBc<int>(g->g32, g->stream, src, bc_host, sigmas_host, labels_host, stats)
                        : RunBc<long long>(g->g64, g->stream, src, bc_host, sigmas_host, labels_host, stats));
}

extern "C" void gunrock_bc_func(struct GunrockGraph *graph_out, const struct GunrockGraph *graph_in,
                                struct GunrockConfig configs, struct GunrockDataType data_type)
{
    // only <int, float, int> exists in the reference (bc_app.cu:160-250)
    if (data_type.VTXID_TYPE != VTXID_INT || data_type.SIZET_TYPE != SIZET_INT || data_type.VALUE_TYPE != VALUE_FLOAT) {
        printf("Not Yet Support This DataType Combination.\n");
        return;
    }
    const long long n = (long long)graph_in->num_nodes, m = (long long)graph_in->num_edges;
    const int *h_off = (const int *)graph_in->row_offsets;
    // bc_app.cu:183-212: -1 (every source) unless a mode picks one
    const int src = grb_pick_source(configs.src_mode, configs.src_node, n, h_off, -1);
    grb_graph_t g = nullptr;
    if (grb_graph_from_host(configs.device, n, m, h_off, 4, (const int32_t *)graph_in->col_indices, nullptr, 0, 4, &g))
        return;
    float *h_bc  = (float *)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1));     // caller owns both (bc_app.cu:70-71,126-128)
    float *h_ebc = (float *)calloc((size_t)(m > 0 ? m : 1), sizeof(float));      // never accumulated by the reference
    grb_stats st;
    memset(&st, 0, sizeof(st));
    const int rc = grb_bc(g, src, h_bc, nullptr, nullptr, &st);
    if (rc == 0) {
        printf("GPU Betweeness Centrality finished in %lf msec.\n", (double)st.elapsed_ms);   // bc_app.cu:130
        graph_out->node_values = h_bc;
        graph_out->edge_values = h_ebc;
    } else {
        free(h_bc); free(h_ebc);
        graph_out->node_values = nullptr;
        graph_out->edge_values = nullptr;
    }
    grb_graph_destroy(g);
    cudaDeviceSynchronize();
}

// ---------------------------------------------------------------------------------------------
extern "C" int grb_topk(int device, int64_t nodes, const int32_t *row_offsets_host, const int32_t *col_offsets_host,
                        int32_t top_nodes, int32_t *node_ids_host, int32_t *in_degrees_host, int32_t *out_degrees_host,
                        grb_stats *stats)
{
    if (grb_select_device(device)) return (int)cudaErrorInvalidDevice;
    if (nodes <= 0 || top_nodes <= 0) return 0;
    if (!row_offsets_host || !col_offsets_host) return (int)cudaErrorInvalidValue;
    const int k = top_nodes > nodes ? (int)nodes : top_nodes;                  // topk_app.cu:175-176
    gunrock::util::DeviceArray<int> out_off, in_off, ids, ind, outd;
    cudaStream_t s = 0;
    GRB_TRY(out_off.FromHost(row_offsets_host, (size_t)nodes + 1, s));
    GRB_TRY(in_off.FromHost(col_offsets_host, (size_t)nodes + 1, s));
    GRB_TRY(ids.Allocate((size_t)k)); GRB_TRY(ind.Allocate((size_t)k)); GRB_TRY(outd.Allocate((size_t)k));
    int launches = 0;
    GRB_TRY(gunrock::app::topk::TopK<int>(out_off.ptr, in_off.ptr, (long long)nodes, k, ids.ptr, ind.ptr, outd.ptr,
                                          &launches, s));
    if (node_ids_host)    GRB_TRY(ids.ToHost(node_ids_host, (size_t)k, s));
    if (in_degrees_host)  GRB_TRY(ind.ToHost(in_degrees_host, (size_t)k, s));
    if (out_degrees_host) GRB_TRY(outd.ToHost(out_degrees_host, (size_t)k, s));
    GRB_TRY(cudaStreamSynchronize(s));
    if (stats) { memset(stats, 0, sizeof(*stats)); stats->gpu_launches = launches; stats->iterations = 1; }
    return 0;
}

extern "C" void gunrock_topk_func(struct GunrockGraph *graph_out, void *node_ids, void *in_degrees, void *out_degrees,
                                  const struct GunrockGraph *graph_in, struct GunrockConfig configs,
                                  struct GunrockDataType data_type)
{
    // only <int, int, int> exists in the reference (topk_app.cu:232-291)
    if (data_type.VTXID_TYPE != VTXID_INT || data_type.SIZET_TYPE != SIZET_INT || data_type.VALUE_TYPE != VALUE_INT) {
        printf("Not Yet Support This DataType Combination.\n");
        return;
    }
    const long long n = (long long)graph_in->num_nodes;
    const int rc = grb_topk(configs.device, n, (const int32_t *)graph_in->row_offsets,
                            (const int32_t *)graph_in->col_offsets, configs.top_nodes,
                            (int32_t *)node_ids, (int32_t *)in_degrees, (int32_t *)out_degrees, nullptr);
    // the reference points graph_out at host vectors that die with the call (build_topk_subgraph,
    // topk_app.cu:78-155): nothing a caller can read; only the counts are meaningful
    if (rc == 0 && graph_out) {
        graph_out->num_nodes = (size_t)(configs.top_nodes > n ? n : (configs.top_nodes > 0 ? configs.top_nodes : 0));
    }
    cudaDeviceSynchronize();
}
