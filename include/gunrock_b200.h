/*
 * gunrock_b200.h -- extended C ABI of libgunrock.so (B200-native build).
 *
 * gunrock/gunrock.h is the reference's interface and is kept as is; it forces
 * a full host->device upload, 32-bit edge offsets and a device->host download
 * on every call (gunrock/app/bfs/bfs_app.cu:146-229).  The entry points below
 * expose the SAME primitives and the two frontier operators with the graph
 * RESIDENT in HBM, 64-bit edge offsets where a graph needs them (R-MAT scale
 * 27), direction-optimising BFS (the reference ships it only as a test
 * program, tests/dobfs/test_dobfs.cu), and the per-level pieces a
 * 1D-partitioned multi-GPU run is assembled from.
 *
 * Conventions: plain C types only; every function returns 0 on success or a
 * cudaError_t value (also reported on stderr, like the reference's GRError);
 * "host" pointers are ordinary host memory, "dev" pointers are device memory
 * on the graph's device.  Nothing here falls back to the CPU.
 *
 * Each declaration cites the reference interface it stands in for.
 */
#ifndef GUNROCK_B200_EXT_H_
#define GUNROCK_B200_EXT_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct grb_graph_s *grb_graph_t;

/* ---- library / device ----------------------------------------------------- */
const char *grb_version(void);
int grb_device_count(void);              /* 0 when no CUDA device is usable */
/* the library keeps released device / pinned buffers for reuse (GRB_POOL=0 disables that);
 * this hands everything cached back to the driver */
void grb_release_cached_memory(void);

/* ---- resident graphs --------------------------------------------------------
 * stands in for ProblemBase::Init's upload (gunrock/app/problem_base.cuh:226-342)
 * and for graphio::BuildRmatGraph / BuildMarketGraph + Csr::FromCoo
 * (gunrock/graphio/rmat.cuh:27-92, market.cuh:57-236, csr.cuh:247-340). */
typedef struct {
    int64_t nodes;          /* global vertex count                               */
    int64_t local_nodes;    /* rows resident on this device                      */
    int64_t edges;          /* CSR entries resident on this device               */
    int32_t offset_bytes;   /* 4 or 8                                            */
    int32_t symmetric;      /* CSR doubles as CSC                                */
    int32_t has_weights;
    int32_t device;
    int32_t log_parts;      /* 1D cyclic partition: 2^log_parts devices          */
    int32_t rank;
} grb_graph_info;

/* upload a host CSR; offset_bytes_in = width of row_offsets (4 or 8);
 * weights may be NULL; symmetric != 0 promises CSR == CSC (enables pull) */
int grb_graph_from_host(int device, int64_t nodes, int64_t edges,
                        const void *row_offsets, int offset_bytes_in,
                        const int32_t *col_indices, const uint32_t *weights,
                        int symmetric, int force_offset_bytes, grb_graph_t *out);

/* One device's slice of a 1D-partitioned graph from host arrays: `local_nodes` rows (row r is global vertex
 * Global(r) of the partition rule in csrc/gunrock/graph/csr.cuh; global_nodes = local_nodes << log_parts), neighbour ids
 * are GLOBAL.  For the grb_mg* contexts of a caller whose partitioned CSR lives on the host; grb_graph_rmat with
 * log_parts > 0 and grb_graph_download produce / return exactly these arrays. */
int grb_graph_slice_from_host(int device, int64_t global_nodes, int log_parts, int rank, int64_t local_nodes,
                              int64_t edges, const void *row_offsets, int offset_bytes_in,
                              const int32_t *col_indices, const uint32_t *weights,
                              int symmetric, int force_offset_bytes, grb_graph_t *out);

/* same, but only the offsets are uploaded: col_indices (and weights) stay in the caller's memory,
 * page-locked and mapped into the device's address space, and the kernels read them over PCIe as
 * they go -- the reference's stream_from_host (gunrock/app/problem_base.cuh:252-276) for graphs
 * larger than device memory or searched once.  The arrays must stay alive and unchanged until
 * grb_graph_destroy.  Memory the caller has already page-locked (cudaHostAlloc / cudaHostRegister)
 * is used as it is; pageable memory is page-locked by this call (for a large array that costs as
 * much as an upload) and released again by grb_graph_destroy of THIS handle -- do not share such
 * arrays between two mapped handles with different lifetimes. */
int grb_graph_from_host_mapped(int device, int64_t nodes, int64_t edges,
                               const void *row_offsets, int offset_bytes_in,
                               const int32_t *col_indices, const uint32_t *weights,
                               int symmetric, int force_offset_bytes, grb_graph_t *out);

/* seeded R-MAT (a,b,c,d + per-level 5% noise exactly as graphio/utils.cuh:84-130),
 * nodes = 2^scale, `edges` generated tuples, undirected => reverse tuples added,
 * then the reference canonicalisation (sort by (row,col), drop duplicates and
 * self loops), all on the device.  With log_parts > 0 only rows v with
 * v % 2^log_parts == rank are kept (1D cyclic partition, problem_base.cuh:186-210).
 * weight_seed != 0 attaches symmetric uint32 weights in [1,64]. */
int grb_graph_rmat(int device, int scale, int64_t edges, uint64_t seed, int undirected,
                   double a, double b, double c, double d,
                   int log_parts, int rank, uint64_t weight_seed,
                   int force_offset_bytes, grb_graph_t *out);

/* build the transpose (CSC) of a directed resident graph, the inverse graph the pull direction of a
 * direction-optimising BFS walks (reference: the inverse Csr handed to DOBFSProblem::Init,
 * app/dobfs/dobfs_problem.cuh:211-260, tests/dobfs/test_dobfs.cu:477-560).  No-op on a symmetric
 * graph; grb_bfs_create does this on demand.  Single-device resident graphs. */
int  grb_graph_build_transpose(grb_graph_t g);
/* build the degree-ordered twin of a resident single-device graph: the same graph with the vertices renumbered by
 * descending degree and every row sorted by the new ids (graph/build.cuh), kept beside the caller's arrays (one more
 * CSR in HBM).  Preprocessing in the sense of the reference's Problem::Init (gunrock/app/problem_base.cuh:279-303:
 * whatever layout the enactor wants is made before the timed Enact).  Searches ask for it with
 * grb_bfs_config.degree_order (built on demand); grb_pagerank sweeps a symmetric graph on its twin when one is
 * attached (ranks come back in the caller's ids; GRB_PR_TWIN=0 opts out). */
int  grb_graph_build_twin(grb_graph_t g);
/* entries of the attached twin (always those of the graph: duplicates and self loops are kept), -1 without a twin */
int  grb_graph_twin_entries(grb_graph_t g, int64_t *entries);
/* symmetric_out = 1 iff the CSR equals its transpose as a multiset of (row, col) entries -- what the
 * `symmetric` promise of grb_graph_from_host means.  Costs a sort of the edge list. */
int  grb_graph_check_symmetric(grb_graph_t g, int *symmetric_out);
int  grb_graph_get_info(grb_graph_t g, grb_graph_info *info);
/* copy the resident CSR back (row_offsets in the graph's offset width) */
int  grb_graph_download(grb_graph_t g, void *row_offsets, int32_t *col_indices, uint32_t *weights);
/* raw device pointers of the resident arrays (for harness-side collectives) */
int  grb_graph_device_ptrs(grb_graph_t g, void **row_offsets, void **col_indices, void **weights);
/* the CUDA stream (cudaStream_t) every operation on this graph is issued on, so a
 * harness can bracket runs with its own events */
int  grb_graph_stream(grb_graph_t g, void **cuda_stream);
void grb_graph_destroy(grb_graph_t g);

/* ---- run statistics ---------------------------------------------------------
 * stands in for the numbers DisplayStats prints (gunrock/app/bfs/bfs_app.cu:63-118) */
typedef struct {
    double   elapsed_ms;        /* CUDA-event time of Enact only (reference bracket, bfs_app.cu:196-200) */
    double   reset_ms;          /* CUDA-event time of Problem::Reset                 */
    int64_t  iterations;        /* search depth / sweeps                             */
    int64_t  edges_visited;     /* sum of out-degrees of visited vertices (bfs_app.cu:82-91) */
    int64_t  nodes_visited;
    uint64_t edges_inspected;   /* neighbour entries the operators actually read     */
    uint64_t total_queued;      /* sum of frontier lengths                           */
    uint64_t algo_bytes;        /* algorithmic HBM bytes, formulas in DESIGN.md      */
    int32_t  push_levels, pull_levels;
    int32_t  gpu_launches;      /* kernels launched inside Enact                     */
    int32_t  reserved;
} grb_stats;

/* per-level trace of the last BFS on a handle (record_levels != 0) */
typedef struct {
    int32_t  level;
    int32_t  kind;              /* 0 push-TWC, 1 pull, 2 push-LB, 3 push-LB (private scan) */
    uint64_t frontier, edges, discovered, candidates;
    float    ms;
    float    switch_ms;         /* part of ms spent converting queue <-> bitmap        */
} grb_level_record;

/* ---- BFS ----------------------------------------------------------------------
 * stands in for gunrock_bfs_func / run_bfs (gunrock/app/bfs/bfs_app.cu:146-229)
 * and for the DOBFS test rig (tests/dobfs/test_dobfs.cu). */
typedef struct {
    int32_t mark_pred;
    int32_t idempotence;
    int32_t traversal_mode;       /* 0 auto, 1 LB, 2 TWC (reference --traversal-mode) */
    int32_t direction_optimized;  /* needs a symmetric graph                          */
    double  alpha, beta;          /* <= 0: library defaults                           */
    double  queue_sizing;         /* <= 0: 1.0                                        */
    int32_t record_levels;
    int32_t host_loop;            /* 0: by size (one cooperative kernel below 2^24 vertices, host-driven
                                     above); 1: one launch set + counter fetch per level, driven by
                                     the host; 2: whole level loop in one cooperative kernel */
    int32_t degree_order;         /* != 0: search the graph's degree-ordered twin (same graph, vertices
                                     renumbered by descending degree, built once per graph handle and kept
                                     next to it: + one CSR of device memory).  Sources, labels and
                                     predecessors stay in the caller's ids.  Resident single-device graphs. */
    int32_t stream_min_edges;     /* host-driven loop: push levels with at least this many frontier edges take the
                                     streaming advance, and filters of idempotent searches with four times as many
                                     items the streaming filter (TMA-staged tiles, hot part of the visited map in
                                     shared memory, claims as reductions + harvest); 0: library default (2^20 on a
                                     degree-ordered twin, never on the caller's vertex order, where the hot range
                                     answers too few probes), 1: never, n > 1: from n entries upward on any order */
} grb_bfs_config;

typedef struct grb_bfs_s *grb_bfs_t;
/* a BFS context owns labels / bitmaps / queues for one graph: create once, run many */
int  grb_bfs_create(grb_graph_t g, const grb_bfs_config *cfg, grb_bfs_t *out);
/* labels_host / preds_host may be NULL (results stay on the device) */
int  grb_bfs_run(grb_bfs_t ctx, int32_t src, int32_t *labels_host, int32_t *preds_host, grb_stats *stats);
/* `count` searches back to back inside the library (no caller code between two searches);
 * stats[i] (optional) receives the i-th search's statistics, labels stay on the device */
int  grb_bfs_run_many(grb_bfs_t ctx, const int32_t *srcs, int32_t count, grb_stats *stats);
int  grb_bfs_device_labels(grb_bfs_t ctx, void **labels_dev);
int  grb_bfs_levels(grb_bfs_t ctx, grb_level_record *out, int32_t cap, int32_t *count);
void grb_bfs_destroy(grb_bfs_t ctx);

/* ---- SSSP / CC / PR -------------------------------------------------------------
 * stand in for gunrock_sssp_func (app/sssp/sssp_app.cu:370), gunrock_cc_func
 * (app/cc/cc_app.cu:265), gunrock_pr_func (app/pr/pr_app.cu:318). */
int grb_sssp(grb_graph_t g, int32_t src, int32_t delta_factor, int32_t mark_pred,
             uint32_t *dist_host, int32_t *preds_host, grb_stats *stats);
int grb_cc(grb_graph_t g, int32_t *labels_host, int64_t *num_components, grb_stats *stats);
/* ranks64_host (fp64, vertex order) and/or sorted (ids, fp32 ranks) outputs; any may be NULL */
int grb_pagerank(grb_graph_t g, int32_t src, double delta, double threshold, int32_t max_iter,
                 double *ranks64_host, int32_t *sorted_ids_host, float *sorted_ranks_host,
                 grb_stats *stats);

/* ---- BC / top-k --------------------------------------------------------------------
 * stand in for gunrock_bc_func (app/bc/bc_app.cu:252) and gunrock_topk_func
 * (app/topk/topk_app.cu:336).  grb_bc: src < 0 accumulates over every source; values are
 * halved as the reference does; any host output may be NULL. */
int grb_bc(grb_graph_t g, int32_t src, float *bc_host, float *sigmas_host, int32_t *labels_host,
           grb_stats *stats);
int grb_topk(int device, int64_t nodes, const int32_t *row_offsets_host, const int32_t *col_offsets_host,
             int32_t top_nodes, int32_t *node_ids_host, int32_t *in_degrees_host, int32_t *out_degrees_host,
             grb_stats *stats);

/* ---- operators, stand-alone (parity tests of the operators themselves) ------------
 * stand in for advance::LaunchKernel (gunrock/oprtr/advance/kernel.cuh:101-129) and
 * filter::Kernel (gunrock/oprtr/filter/kernel.cuh:738-757) run with a probe functor:
 *   CondEdge(s,d,e)  : edge_hits[e] += 1; src_sum[d] += s; return ((s + d) % 3) != 0
 *   ApplyEdge(s,d,e) : dst_hits[d] += 1
 *   CondFilter(v)    : return v % 2 == 0        ApplyFilter(v): dst_hits[v] += 1
 * mode: 0 TWC, 1 LB, plus 16 * frontier type (0 V2V, 1 V2E, 2 E2V, 3 E2E: an input edge id stands
 * for the vertex that edge points to, an output edge frontier receives CSR entry indices --
 * reference advance::TYPE, oprtr/advance/kernel_policy.cuh:54-60).
 * output: 0 none, 1 compact, 2 tombstone (LB only).
 * All array arguments are host pointers; out_queue must hold out_cap ids. */
int grb_test_advance(grb_graph_t g, int mode, int output,
                     const int32_t *in_queue, int64_t in_len,
                     int32_t *out_queue, int64_t out_cap, int64_t *out_len,
                     int32_t *edge_hits, int64_t *src_sum, int32_t *dst_hits,
                     uint64_t *edges_inspected);
/* neighbour-reduce advance (reference: advance::LaunchKernel with R_OP / R_TYPE,
 * oprtr/advance/kernel.cuh:731-761): reduced[i] = op over the edges (s, d, e) of input item i that
 * pass the probe functor's CondEdge of values[d] (rtype 1, VERTEX: n values) or values[e] (rtype 2,
 * EDGE: m values); identity when none passes.  op: the reference's REDUCE_OP numbering
 * (1 PLUS, 3 MULTIPLIES, 5 BIT_OR, 6 BIT_AND, 7 BIT_XOR, 8 MAXIMUM, 9 MINIMUM).  atype: frontier
 * type as above.  out_queue (optional) receives the compacted output frontier. */
int grb_test_neighbor_reduce(grb_graph_t g, int op, int rtype, int atype,
                             const int32_t *in_queue, int64_t in_len, const int32_t *values,
                             int32_t *reduced, int32_t *out_queue, int64_t out_cap, int64_t *out_len);
int grb_test_filter(int device, int64_t nodes, int compact,
                    const int32_t *in_queue, int64_t in_len,
                    int32_t *out_queue, int64_t out_cap, int64_t *out_len,
                    int32_t *dst_hits);
/* The near/far split of the priority queue on its own (priority_queue::Bisect; replaces the reference's
 * gunrock/priority_queue/kernel.cuh:404-503 Bisect / RemoveInvalid): items whose score (scores[item]) is below
 * `lower` are dropped, those in [lower, upper) go to near_out, the others to far_out unless their bit in far_mark
 * is already set (in_is_far = 0) -- or, when the input IS a far pile (in_is_far = 1), leaving items clear their bit.
 * far_mark is read and written back; min_far_score receives the smallest score sent to the far pile (~0 if none). */
int grb_test_bisect(int device, int64_t nodes, const uint32_t *scores, const int32_t *in_items, int64_t in_len,
                    int in_is_far, uint32_t lower, uint32_t upper, uint32_t *far_mark,
                    int32_t *near_out, int64_t near_cap, int64_t *near_len,
                    int32_t *far_out, int64_t far_cap, int64_t *far_len, uint32_t *min_far_score);

/* The row bins of the fixed-row-set passes on their own (oprtr/row_bins.cuh; the passes replace the reference's
 * load-balanced advance, gunrock/oprtr/edge_map_partitioned/kernel.cuh:242-559, where the rows are known in advance:
 * PageRank's sweeps, gunrock/app/pr/pr_enactor.cuh:346-487, and CC's hook pass, gunrock/app/cc/cc_enactor.cuh:443-560).
 * queue: qlen global vertex ids in host memory, or NULL for rows 0 .. qlen.  Rows of 1..16 entries go to small_rows,
 * 17..2048 to medium_rows, longer ones to big_rows and, cut into 2048-entry pieces, to big_pieces ((row, piece) pairs,
 * room for piece_cap of them); each of the three row arrays needs room for qlen ids.  counts[8] = rows per list,
 * pieces, entries of all listed rows, and the CTAs a pass assigns to pieces / medium rows / small rows. */
int grb_test_row_bins(grb_graph_t g, const int32_t *queue, int64_t qlen, int32_t *small_rows, int32_t *medium_rows,
                      int32_t *big_rows, int32_t *big_pieces, int64_t piece_cap, int64_t *counts);

/* ---- the LB advance for callers that cannot allocate --------------------------------
 * No counterpart in the reference: this entry point was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  It is grb_test_advance(mode = 1 (LB), V2V,
 * output 0 none / 1 compact) with the same probe functor, except that
 *   - EVERY array argument is a DEVICE pointer owned by the caller (the graph handle already
 *     holds the CSR on its device), the kernels are queued on `cuda_stream` (a cudaStream_t,
 *     NULL = the default stream) and the call returns without synchronising;
 *   - edge_hits_dev [edges], src_sum_dev [nodes], dst_hits_dev [nodes] are ACCUMULATED into: the
 *     caller zeroes them; out_queue_dev holds out_cap ids;
 *   - counters_dev (optional) receives eight 64-bit words: out_len, out_degree (not tracked: 0),
 *     edges inspected, degree sum of the input frontier, aux0, aux1, overflow (!= 0: out_cap or
 *     GRB_NOALLOC_MAX_EDGES exceeded, the output is incomplete), reserved;
 *   - it allocates NOTHING (no cudaMalloc / cudaFree / cudaMallocHost, none of the library's pooled
 *     arrays): the scratch of the LB mode lives in a static __device__ array of
 *     GRB_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until exit.
 * Limits: in_len <= GRB_NOALLOC_MAX_ITEMS (a longer frontier is rejected with
 * cudaErrorInvalidValue and a message on stderr, nothing is launched); at most
 * GRB_NOALLOC_MAX_EDGES neighbour entries per call (known on the device only: reported through
 * the overflow counter); a whole graph, not a slice of a partition.
 * Every call on a device shares the one scratch array: calls that overlap on the GPU (two streams
 * with no ordering between them, two host threads) race on it -- serialise them. */
#define GRB_NOALLOC_MAX_ITEMS     262144        /* 2^18 frontier items per call */
#define GRB_NOALLOC_MAX_EDGES     16777216      /* 2^24 neighbour entries per call */
#define GRB_NOALLOC_SCRATCH_BYTES 4328960       /* static device memory held per GPU */
int grb_advance_lb_noalloc(grb_graph_t g, int output,
                           const int32_t *in_queue_dev, int64_t in_len,
                           int32_t *out_queue_dev, int64_t out_cap,
                           int32_t *edge_hits_dev, int64_t *src_sum_dev, int32_t *dst_hits_dev,
                           uint64_t *counters_dev, void *cuda_stream);

/* ---- 1D-partitioned BFS, one device per process -----------------------------------
 * No reference counterpart (the reference's multi-GPU path is a TODO,
 * gunrock/app/problem_base.cuh:336-338); the striping rule is the reference's
 * (:186-210).  The harness owns the communicator (torch.distributed / NCCL) and runs
 * its collectives on the context's stream between the steps below; every step is
 * asynchronous on that stream.
 *
 * All bitmaps are GLOBAL, one bit per vertex, laid out owner-major: device r's bits
 * are words [r * seg_words, (r + 1) * seg_words), seg_words = local_nodes / 32.
 *   push level : grb_mgbfs_push -> all-to-all(next map -> recv map, seg_words words per
 *                peer) -> grb_mgbfs_absorb
 *   pull level : all-gather(frontier map `cur`, own segment in place) -> grb_mgbfs_pull
 * After either, the 64-byte counter block (buffer 6) holds, as 8 uint64:
 *   [0] vertices this device added to the next frontier   [1] their out-degree sum
 *   [2] neighbour entries inspected  [4] pull: unvisited candidates  [6] overflow flag  */
typedef struct grb_mgbfs_s *grb_mgbfs_t;
typedef struct {
    uint64_t frontier_local;     /* owned vertices in the next frontier            */
    uint64_t frontier_edges;     /* their out-degree sum                           */
    uint64_t edges_inspected;
    uint64_t candidates;
    uint64_t overflow;
    uint64_t gpu_launches;       /* kernels launched by this context so far        */
} grb_mg_level;

int  grb_mgbfs_create(grb_graph_t g, grb_mgbfs_t *out);
/* labels <- -1, bitmaps <- 0, source marked visited everywhere, queued on its owner;
 * src_degree is the source's out-degree on the owner, 0 elsewhere (synchronises) */
int  grb_mgbfs_reset(grb_mgbfs_t ctx, int32_t src, uint64_t *src_degree);
/* device pointers the harness hands to the collectives:
 * which: 0 visited map, 1 frontier map 0, 2 frontier map 1, 3 next map, 4 recv map,
 *        5 labels (int32, local rows), 6 counter block */
int  grb_mgbfs_buffer(grb_mgbfs_t ctx, int which, void **dev_ptr, uint64_t *bytes);
/* push level, part 1: clears the counters and the next map, expands the local queue of
 * `queue_len` vertices; every unvisited neighbour sets its bit in the next map */
int  grb_mgbfs_push(grb_mgbfs_t ctx, int32_t level, uint64_t queue_len, uint64_t edges_hint);
/* push level, part 2 (after the all-to-all): OR of the received segments & ~visited ->
 * labels level+1, own segment of frontier map `cur_out`, new local queue, counters */
int  grb_mgbfs_absorb(grb_mgbfs_t ctx, int32_t level, int cur_out);
/* pull level: clears the counters, reads frontier map `cur` (all segments), writes the
 * OWN segment of map cur^1, labels level+1, updates the own visited segment */
int  grb_mgbfs_pull(grb_mgbfs_t ctx, int32_t level, int cur);
/* visited |= frontier map `cur`, all segments (refreshes the remote cull hints) */
int  grb_mgbfs_merge_visited(grb_mgbfs_t ctx, int cur);
/* local queue -> own segment of map `cur` (map cleared first); and back */
int  grb_mgbfs_queue_to_map(grb_mgbfs_t ctx, int cur, uint64_t queue_len);
int  grb_mgbfs_map_to_queue(grb_mgbfs_t ctx, int cur);
/* host copy of this device's counter block (synchronises the stream) */
int  grb_mgbfs_counters(grb_mgbfs_t ctx, grb_mg_level *out);
int  grb_mgbfs_sync(grb_mgbfs_t ctx);
int  grb_mgbfs_stream(grb_mgbfs_t ctx, void **cuda_stream);
int  grb_mgbfs_labels(grb_mgbfs_t ctx, int32_t *labels_host);
void grb_mgbfs_destroy(grb_mgbfs_t ctx);

/* ---- peer-memory exchange: the same levels with NVLink stores instead of collectives ----
 * Each process exports CUDA IPC handles of the buffers its peers write (frontier map 0,
 * frontier map 1, receive map, mailbox), the harness moves the handle bytes between the
 * processes (any transport; torch.distributed all_gather in gunrockinst_b200/partitioned.py),
 * and every process imports all of them.  After that grb_mgbfs_run_peer runs a whole
 * direction-optimising search natively: pull sweeps publish their frontier words on every
 * device as they are produced, push levels store the lines of the discovery map that hold a
 * bit straight into the owner's receive buffer, and level counters / barriers are flag words
 * in the peers' mailboxes -- no collective library call inside a search.  All ranks must call
 * it with the same arguments; a rank that waits more than 4 s for a peer gives up with
 * cudaErrorLaunchTimeout.  With one device no import is needed. */
#define GRB_IPC_HANDLE_BYTES 64
#define GRB_MG_PEER_BUFFERS  4
typedef struct {
    double   elapsed_ms;         /* CUDA events on the context's stream, reset included */
    int64_t  iterations, push_levels, pull_levels, gpu_launches;
    uint64_t edges_visited;      /* sum of out-degrees of the visited vertices, all devices */
    uint64_t edges_inspected;    /* neighbour entries inspected, all devices              */
    uint64_t exchange_bytes;     /* bytes this device stored into peer memory             */
    uint64_t nodes_visited;
} grb_mg_stats;
typedef struct {
    int32_t  level, kind;        /* kind: 0 push, 1 pull */
    uint64_t frontier, discovered, edges, max_edges, exchange_bytes;
    uint64_t candidates;         /* pull: unvisited vertices examined, all devices */
    float    compute_ms, exchange_ms;
} grb_mg_trace;
int  grb_mgbfs_peer_export(grb_mgbfs_t ctx, unsigned char *handles /* [GRB_MG_PEER_BUFFERS * 64] */);
int  grb_mgbfs_peer_import(grb_mgbfs_t ctx, int world,
                           const unsigned char *all_handles /* [world][GRB_MG_PEER_BUFFERS * 64] */,
                           uint64_t global_edges);
int  grb_mgbfs_run_peer(grb_mgbfs_t ctx, int32_t src, int direction_optimized, double alpha, double beta,
                        int record_levels, grb_mg_stats *out);
int  grb_mgbfs_peer_levels(grb_mgbfs_t ctx, grb_mg_trace *out, int32_t cap, int32_t *count);

/* ---- 1D-partitioned SSSP over peer memory -----------------------------------------------
 * Same partition and handle hand-over as the BFS above (3 buffers per device: the 64-bit
 * distance|predecessor words, the touched bitmap, the mailbox).  A remote relaxation is a 64-bit
 * atomicMin issued over NVLink by the advance kernel that found the shorter path; the owner
 * collects what was touched into its next frontier.  Relaxation arithmetic and outputs as the
 * reference's SSSP (gunrock/app/sssp/sssp_functor.cuh:52-66, sssp_problem.cuh:325,372): uint32
 * distances, UINT_MAX = unreachable, predecessor = own id for the source and unreached vertices.
 * The graph needs edge weights.  grb_mg_stats: iterations, gpu_launches, edges_inspected =
 * relaxations executed on all devices, nodes_visited = vertices queued over all iterations. */
#define GRB_MGSSSP_PEER_BUFFERS 3
typedef struct grb_mgsssp_s *grb_mgsssp_t;
int  grb_mgsssp_create(grb_graph_t g, grb_mgsssp_t *out);
int  grb_mgsssp_peer_export(grb_mgsssp_t ctx, unsigned char *handles /* [GRB_MGSSSP_PEER_BUFFERS * 64] */);
int  grb_mgsssp_peer_import(grb_mgsssp_t ctx, int world,
                            const unsigned char *all_handles /* [world][GRB_MGSSSP_PEER_BUFFERS * 64] */);
int  grb_mgsssp_run(grb_mgsssp_t ctx, int32_t src, grb_mg_stats *out);     /* every rank, same src */
/* distances / predecessors of the rows this device owns, local-row order (either may be NULL) */
int  grb_mgsssp_result(grb_mgsssp_t ctx, uint32_t *dist_host, int32_t *preds_host);
void grb_mgsssp_destroy(grb_mgsssp_t ctx);

/* ---- 1D-partitioned PageRank over peer memory -------------------------------------------
 * Same partition and handle hand-over (3 buffers per device: the two generations of the
 * replicated contributions rank / degree, the two generations of the replicated "degree
 * reached zero" map of the pruning phase, the mailbox).  A sweep gathers along the in-edges
 * of the owned rows from the local replica (no remote atomics); the kernel that finishes the
 * sweep stores every owned vertex's new contribution into all replicas over NVLink.  Pruning,
 * damping, base term and the moving test are the reference's (gunrock/app/pr/
 * pr_functor.cuh:32-175, pr_enactor.cuh:247-329,346-487), in fp64 like grb_pagerank.  The
 * graph must be symmetric.  grb_mg_stats: iterations = sweeps, push_levels = pruning sweeps,
 * nodes_visited = vertices that survived the pruning (all devices), edges_inspected = entries
 * walked on all devices, exchange_bytes = bytes this device stored into peer memory,
 * edges_visited = microseconds of the power iteration alone. */
#define GRB_MGPR_PEER_BUFFERS 3
typedef struct grb_mgpr_s *grb_mgpr_t;
int  grb_mgpr_create(grb_graph_t g, grb_mgpr_t *out);
int  grb_mgpr_peer_export(grb_mgpr_t ctx, unsigned char *handles /* [GRB_MGPR_PEER_BUFFERS * 64] */);
int  grb_mgpr_peer_import(grb_mgpr_t ctx, int world,
                          const unsigned char *all_handles /* [world][GRB_MGPR_PEER_BUFFERS * 64] */);
/* every rank, same arguments (src < 0: global PageRank, else personalised) */
int  grb_mgpr_run(grb_mgpr_t ctx, int32_t src, double delta, double threshold, int32_t max_iter, grb_mg_stats *out);
/* fp64 ranks of the rows this device owns, local-row order */
int  grb_mgpr_result(grb_mgpr_t ctx, double *ranks_host);
void grb_mgpr_destroy(grb_mgpr_t ctx);

/* ---- 1D-partitioned connected components over peer memory ------------------------------
 * Same partition, handle hand-over and buffers as the partitioned SSSP.  Minimum-label
 * propagation: every vertex starts with its own id in the frontier, a smaller label travels along
 * an edge as a 64-bit atomicMin into the owner's memory (over NVLink when remote), the owner
 * re-queues rows whose label dropped; ends when no device holds an unsent improvement.  Labels as
 * grb_cc / gunrock_cc_func: the smallest vertex id of the component (gunrock/app/cc/
 * cc_problem.cuh:164-170 after canonicalisation).  The graph must be symmetric.  grb_mg_stats:
 * iterations, gpu_launches, edges_inspected = label offers examined on all devices,
 * nodes_visited = vertices queued over all iterations. */
#define GRB_MGCC_PEER_BUFFERS 3
typedef struct grb_mgcc_s *grb_mgcc_t;
int  grb_mgcc_create(grb_graph_t g, grb_mgcc_t *out);
int  grb_mgcc_peer_export(grb_mgcc_t ctx, unsigned char *handles /* [GRB_MGCC_PEER_BUFFERS * 64] */);
int  grb_mgcc_peer_import(grb_mgcc_t ctx, int world,
                          const unsigned char *all_handles /* [world][GRB_MGCC_PEER_BUFFERS * 64] */);
int  grb_mgcc_run(grb_mgcc_t ctx, grb_mg_stats *out);                      /* every rank */
/* labels of the rows this device owns, local-row order (may be NULL); *local_roots = owned vertices
 * that are the label of their component (their sum over the devices is the component count) */
int  grb_mgcc_result(grb_mgcc_t ctx, int32_t *labels_host, int64_t *local_roots);
void grb_mgcc_destroy(grb_mgcc_t ctx);

#ifdef __cplusplus
}
#endif
#endif /* GUNROCK_B200_EXT_H_ */
