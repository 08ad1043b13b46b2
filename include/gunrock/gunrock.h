/*
 * gunrock.h -- C ABI of libgunrock (B200-native build).
 *
 * This header is the drop-in boundary: it declares, with identical names,
 * argument order, struct layouts and enumerator values, the C interface that
 * the reference library publishes in /root/reference/gunrock/gunrock.h
 * (enums :24-47, GunrockDataType :52-57, GunrockGraph :62-72, SrcMode :77-82,
 * GunrockConfig :87-100, entry points :106-151).  A C program written against
 * the reference header (e.g. shared_lib_tests/test_bfs.c, test_cc.c,
 * test_sssp.c, test_pr.c) compiles against this file and links against this
 * repository's libgunrock.so without modification.
 *
 * Only layout-relevant content is restated here; all behaviour behind the
 * entry points is implemented from scratch as sm_100a CUDA
 * (gunrockinst_b200/csrc).  There is no CPU fallback: every entry point
 * prints a diagnostic to stderr and returns without producing a result if no
 * CUDA device can be initialised.
 *
 * Ownership rules (same as the reference, see bfs_app.cu:165,211):
 *   - input arrays are borrowed and never modified;
 *   - graph_out->node_values is allocated by the library with malloc() and
 *     must be released by the caller with free();
 *   - predecessor / node_ids / page_rank buffers are caller-allocated.
 */
#ifndef GUNROCK_B200_GUNROCK_H_
#define GUNROCK_B200_GUNROCK_H_

#include <stdlib.h>
#include <stdbool.h>

/* ---- element-type selectors --------------------------------------------- */

/* type of a vertex identifier */
enum VertexIdType {
    VTXID_INT                       /* 32-bit signed */
};

/* type of an edge offset / graph size */
enum SizeTType {
    SIZET_INT                       /* 32-bit signed */
};

/* type of the per-vertex / per-edge value a primitive works on */
enum ValueType {
    VALUE_INT,                      /* BFS depth, CC label           */
    VALUE_UINT,                     /* SSSP distance / edge weight   */
    VALUE_FLOAT                     /* PageRank score                */
};

/* the <VertexId, SizeT, Value> triple of one call */
struct GunrockDataType {
    enum VertexIdType VTXID_TYPE;
    enum SizeTType    SIZET_TYPE;
    enum ValueType    VALUE_TYPE;
};

/* ---- graph container ------------------------------------------------------ */

/*
 * CSR (and optionally CSC) view of a graph plus result slots.  Callers fill
 * num_nodes, num_edges, row_offsets, col_indices (and edge_values for SSSP);
 * every other field may hold garbage on input and is never read.
 */
struct GunrockGraph {
    size_t  num_nodes;              /* |V|                                    */
    size_t  num_edges;              /* |E| (directed CSR entries)             */
    void   *row_offsets;            /* CSR: num_nodes+1 offsets               */
    void   *col_indices;            /* CSR: num_edges neighbour ids           */
    void   *col_offsets;            /* CSC offsets (unused by this build)     */
    void   *row_indices;            /* CSC indices (unused by this build)     */
    void   *node_values;            /* OUT: per-vertex result, malloc()ed     */
    void   *edge_values;            /* IN: per-edge weights (SSSP)            */
};

/* ---- run configuration ---------------------------------------------------- */

/* how the traversal source is chosen */
enum SrcMode {
    manually,                       /* use GunrockConfig.src_node             */
    randomize,                      /* draw from libc rand() like the ref.    */
    largest_degree                  /* first vertex of maximum out-degree     */
};

/* passed BY VALUE to every entry point; irrelevant fields may be garbage */
struct GunrockConfig {
    bool  mark_pred;                /* BFS/SSSP: record predecessors          */
    bool  idempotence;              /* BFS: idempotent (duplicate-tolerant)   */
    int   src_node;                 /* source when src_mode == manually       */
    int   device;                   /* CUDA device ordinal                    */
    int   max_iter;                 /* PR/BC: iteration cap                   */
    int   top_nodes;                /* PR/topk: number of results wanted      */
    int   delta_factor;             /* SSSP: delta-stepping bucket factor     */
    float delta;                    /* PR: damping factor                     */
    float error;                    /* PR: convergence threshold              */
    float queue_size;               /* frontier queue sizing factor           */
    enum  SrcMode src_mode;
};

#ifdef __cplusplus
extern "C" {
#endif

/* breadth-first search: node_values = int depth, -1 if unreachable */
void gunrock_bfs_func(
    struct GunrockGraph       *graph_out,
    const struct GunrockGraph *graph_in,
    struct GunrockConfig       configs,
    struct GunrockDataType     data_type);

/* betweenness centrality: node_values = malloc'ed float[n] (x0.5, summed over every source when
 * src_node == -1), edge_values = malloc'ed float[m] of zeros (the reference never accumulates it) */
void gunrock_bc_func(
    struct GunrockGraph       *graph_out,
    const struct GunrockGraph *graph_in,
    struct GunrockConfig       configs,
    struct GunrockDataType     data_type);

/* connected components: node_values = int label = smallest vertex id */
void gunrock_cc_func(
    struct GunrockGraph       *graph_out,
    const struct GunrockGraph *graph_in,
    struct GunrockConfig       configs,
    struct GunrockDataType     data_type);

/* single-source shortest paths: node_values = unsigned distance
 * (UINT_MAX if unreachable); predecessor[] receives num_nodes ints */
void gunrock_sssp_func(
    struct GunrockGraph       *graph_out,
    void                      *predecessor,
    const struct GunrockGraph *graph_in,
    struct GunrockConfig       configs,
    struct GunrockDataType     data_type);

/* PageRank: node_ids / page_rank receive the vertices in descending rank */
void gunrock_pr_func(
    struct GunrockGraph       *graph_out,
    void                      *node_ids,
    void                      *page_rank,
    const struct GunrockGraph *graph_in,
    struct GunrockConfig       configs,
    struct GunrockDataType     data_type);

/* degree-centrality top-k: caller-allocated int[top_nodes] outputs, ordered by in+out degree
 * (descending, ties by ascending vertex id); graph_in needs col_offsets as well */
void gunrock_topk_func(
    struct GunrockGraph       *graph_out,
    void                      *node_ids,
    void                      *in_degrees,
    void                      *out_degrees,
    const struct GunrockGraph *graph_in,
    struct GunrockConfig       configs,
    struct GunrockDataType     data_type);

#ifdef __cplusplus
}
#endif

#endif /* GUNROCK_B200_GUNROCK_H_ */
