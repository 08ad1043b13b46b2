// grb_graph.cu -- resident graph handles of the extended C ABI (include/gunrock_b200.h).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"

extern "C" const char *grb_version(void) { return "gunrockinst_b200 0.1 (sm_100a)"; }

extern "C" void grb_release_cached_memory(void)
{
    gunrock::util::BlockCache::Device().Trim();
    gunrock::util::BlockCache::Pinned().Trim();
}

extern "C" int grb_device_count(void)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return 0; }
    return count;
}

static int GraphFromHost(int device, int64_t nodes, int64_t edges,
                         const void *row_offsets, int offset_bytes_in,
                         const int32_t *col_indices, const uint32_t *weights,
                         int symmetric, int force_offset_bytes, bool mapped, grb_graph_t *out)
{
    if (!out || !row_offsets || (edges > 0 && !col_indices) || (offset_bytes_in != 4 && offset_bytes_in != 8))
        return (int)cudaErrorInvalidValue;
    *out = nullptr;
    int rc = grb_select_device(device);
    if (rc) return rc;
    int ob = force_offset_bytes ? force_offset_bytes : (edges >= (int64_t)0x7fffffff ? 8 : 4);
    if (ob != 4 && ob != 8) return (int)cudaErrorInvalidValue;
    if (ob == 4 && edges >= (int64_t)0x7fffffff) return (int)cudaErrorInvalidValue;
    grb_graph_s *h = new grb_graph_s();
    h->offset_bytes = ob;
    cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) {
        if (ob == 4) {
            h->g32 = new Csr32(); h->g32->device = device; h->g32->symmetric = symmetric != 0;
            e = (offset_bytes_in == 4)
                ? h->g32->FromHost(nodes, edges, (const int *)row_offsets, col_indices, weights, h->stream, mapped)
                : h->g32->FromHost(nodes, edges, (const long long *)row_offsets, col_indices, weights, h->stream, mapped);
        } else {
            h->g64 = new Csr64(); h->g64->device = device; h->g64->symmetric = symmetric != 0;
            e = (offset_bytes_in == 4)
                ? h->g64->FromHost(nodes, edges, (const int *)row_offsets, col_indices, weights, h->stream, mapped)
                : h->g64->FromHost(nodes, edges, (const long long *)row_offsets, col_indices, weights, h->stream, mapped);
        }
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) { gunrock::util::GRError(e, "grb_graph_from_host", __FILE__, __LINE__); delete h; return (int)e; }
    *out = h;
    return 0;
}

extern "C" int grb_graph_from_host(int device, int64_t nodes, int64_t edges,
                                   const void *row_offsets, int offset_bytes_in,
                                   const int32_t *col_indices, const uint32_t *weights,
                                   int symmetric, int force_offset_bytes, grb_graph_t *out)
{
    return GraphFromHost(device, nodes, edges, row_offsets, offset_bytes_in, col_indices, weights, symmetric,
                         force_offset_bytes, false, out);
}

// one device's slice of a 1D-partitioned graph, from host arrays (local rows, GLOBAL neighbour ids): what a caller who
// keeps the partitioned CSR on the host uploads per rank (the reference declares the striping, problem_base.cuh:186-210,
// and never builds the slices)
extern "C" int grb_graph_slice_from_host(int device, int64_t global_nodes, int log_parts, int rank, int64_t local_nodes,
                                         int64_t edges, const void *row_offsets, int offset_bytes_in,
                                         const int32_t *col_indices, const uint32_t *weights,
                                         int symmetric, int force_offset_bytes, grb_graph_t *out)
{
    if (log_parts < 0 || log_parts > 3 || rank < 0 || rank >= (1 << log_parts) || local_nodes < 0 ||
        (local_nodes << log_parts) != global_nodes)
        return (int)cudaErrorInvalidValue;
    int rc = GraphFromHost(device, local_nodes, edges, row_offsets, offset_bytes_in, col_indices, weights, symmetric,
                           force_offset_bytes, false, out);
    if (rc) return rc;
    grb_graph_s *h = *out;
    if (h->g32) { h->g32->nodes = global_nodes; h->g32->part.log_parts = log_parts; h->g32->part.rank = rank; }
    else        { h->g64->nodes = global_nodes; h->g64->part.log_parts = log_parts; h->g64->part.rank = rank; }
    return 0;
}

// the reference's stream_from_host (gunrock/app/problem_base.cuh:252-276)
extern "C" int grb_graph_from_host_mapped(int device, int64_t nodes, int64_t edges,
                                          const void *row_offsets, int offset_bytes_in,
                                          const int32_t *col_indices, const uint32_t *weights,
                                          int symmetric, int force_offset_bytes, grb_graph_t *out)
{
    return GraphFromHost(device, nodes, edges, row_offsets, offset_bytes_in, col_indices, weights, symmetric,
                         force_offset_bytes, true, out);
}

extern "C" int grb_graph_build_transpose(grb_graph_t g)
{
    if (!g) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    GRB_TRY(g->EnsureTranspose());
    return 0;
}

extern "C" int grb_graph_build_twin(grb_graph_t g)
{
    if (!g) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    GRB_TRY(g->EnsureTwin());
    GRB_TRY(cudaStreamSynchronize(g->stream));
    return 0;
}

extern "C" int grb_graph_twin_entries(grb_graph_t g, int64_t *entries)
{
    if (!g || !entries) return (int)cudaErrorInvalidValue;
    *entries = g->twin32 ? (int64_t)g->twin32->edges : (g->twin64 ? (int64_t)g->twin64->edges : -1);
    return 0;
}

// CSR == its transpose?  (graph/build.cuh: both are canonical sorted CSRs of the same multiset iff equal)
template <typename G>
static cudaError_t CheckSymmetric(G &g, cudaStream_t s, int *out)
{
    using namespace gunrock;
    *out = 0;
    if (g.part.log_parts != 0 || g.streams_from_host)
        return util::GRError(cudaErrorInvalidValue, "symmetry check: resident single-device graphs only", __FILE__, __LINE__);
    G t;
    t.nodes = t.local_nodes = g.nodes; t.edges = g.edges; t.device = g.device;
    // transpose into a scratch object: borrow the forward arrays, keep the result in t.csc_*
    t.row_offsets.Borrow(g.row_offsets.ptr, g.row_offsets.size);
    t.col_indices.Borrow(g.col_indices.ptr, g.col_indices.size);
    GR_TRY(graph::BuildTranspose(t, s));
    util::DeviceArray<unsigned long long> diff;
    GR_TRY(diff.Allocate(1));
    GR_TRY(diff.Zero(s));
    // the forward rows must be sorted for an element-wise comparison: sort a copy through the same path
    G f;
    f.nodes = f.local_nodes = g.nodes; f.edges = g.edges; f.device = g.device;
    f.row_offsets.Borrow(t.csc_offsets.ptr, t.csc_offsets.size);
    f.col_indices.Borrow(t.csc_indices.ptr, t.csc_indices.size);
    GR_TRY(graph::BuildTranspose(f, s));        // transpose of the transpose = the forward graph, rows sorted
    graph::CountDifferencesKernel<<<util::FillGrid((unsigned long long)g.edges + (unsigned long long)g.nodes + 1), 256, 0, s>>>(
        f.csc_offsets.ptr, t.csc_offsets.ptr, (unsigned long long)g.nodes + 1,
        f.csc_indices.ptr, t.csc_indices.ptr, (unsigned long long)g.edges, diff.ptr);
    GR_TRY(cudaGetLastError());
    unsigned long long h = 1;
    GR_TRY(cudaMemcpyAsync(&h, diff.ptr, sizeof(h), cudaMemcpyDeviceToHost, s));
    GR_TRY(cudaStreamSynchronize(s));
    *out = h == 0 ? 1 : 0;
    return cudaSuccess;
}

extern "C" int grb_graph_check_symmetric(grb_graph_t g, int *symmetric_out)
{
    if (!g || !symmetric_out) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    if (g->g32) GRB_TRY(CheckSymmetric(*g->g32, g->stream, symmetric_out));
    else        GRB_TRY(CheckSymmetric(*g->g64, g->stream, symmetric_out));
    return 0;
}

extern "C" int grb_graph_get_info(grb_graph_t g, grb_graph_info *info)
{
    if (!g || !info) return (int)cudaErrorInvalidValue;
    memset(info, 0, sizeof(*info));
    info->nodes = g->nodes(); info->local_nodes = g->local_nodes(); info->edges = g->edges();
    info->offset_bytes = g->offset_bytes; info->device = g->device();
    if (g->g32) {
        info->symmetric = g->g32->symmetric; info->has_weights = g->g32->edge_values.size != 0;
        info->log_parts = g->g32->part.log_parts; info->rank = g->g32->part.rank;
    } else {
        info->symmetric = g->g64->symmetric; info->has_weights = g->g64->edge_values.size != 0;
        info->log_parts = g->g64->part.log_parts; info->rank = g->g64->part.rank;
    }
    return 0;
}

extern "C" int grb_graph_download(grb_graph_t g, void *row_offsets, int32_t *col_indices, uint32_t *weights)
{
    if (!g) return (int)cudaErrorInvalidValue;
    GRB_TRY(cudaSetDevice(g->device()));
    const size_t ln = (size_t)g->local_nodes(), m = (size_t)g->edges();
    if (g->g32) {
        if (row_offsets) GRB_TRY(g->g32->row_offsets.ToHost((int *)row_offsets, ln + 1, g->stream));
        if (col_indices) GRB_TRY(g->g32->col_indices.ToHost(col_indices, m, g->stream));
        if (weights && g->g32->edge_values.size) GRB_TRY(g->g32->edge_values.ToHost(weights, m, g->stream));
    } else {
        if (row_offsets) GRB_TRY(g->g64->row_offsets.ToHost((long long *)row_offsets, ln + 1, g->stream));
        if (col_indices) GRB_TRY(g->g64->col_indices.ToHost(col_indices, m, g->stream));
        if (weights && g->g64->edge_values.size) GRB_TRY(g->g64->edge_values.ToHost(weights, m, g->stream));
    }
    GRB_TRY(cudaStreamSynchronize(g->stream));
    return 0;
}

extern "C" int grb_graph_device_ptrs(grb_graph_t g, void **row_offsets, void **col_indices, void **weights)
{
    if (!g) return (int)cudaErrorInvalidValue;
    if (g->g32) {
        if (row_offsets) *row_offsets = g->g32->row_offsets.ptr;
        if (col_indices) *col_indices = g->g32->col_indices.ptr;
        if (weights) *weights = g->g32->edge_values.size ? g->g32->edge_values.ptr : nullptr;
    } else {
        if (row_offsets) *row_offsets = g->g64->row_offsets.ptr;
        if (col_indices) *col_indices = g->g64->col_indices.ptr;
        if (weights) *weights = g->g64->edge_values.size ? g->g64->edge_values.ptr : nullptr;
    }
    return 0;
}

extern "C" int grb_graph_stream(grb_graph_t g, void **cuda_stream)
{
    if (!g || !cuda_stream) return (int)cudaErrorInvalidValue;
    *cuda_stream = (void *)g->stream;
    return 0;
}

extern "C" void grb_graph_destroy(grb_graph_t g)
{
    if (!g) return;
    cudaSetDevice(g->device());
    delete g;
}
