// mgbfs_capi.cu -- C entry points of the 1D-partitioned BFS steps (include/gunrock_b200.h,
// grb_mgbfs_*).  No reference counterpart: gunrock/app/problem_base.cuh:336-338 leaves the
// multi-GPU path as a TODO; the striping rule it declares (:186-210) is the one used here.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/app/bfs/mgbfs_enactor.cuh"

struct MgRunnerBase {
    virtual ~MgRunnerBase() {}
    virtual cudaError_t Reset(int src, unsigned long long *src_degree) = 0;
    virtual cudaError_t Buffer(int which, void **ptr, unsigned long long *bytes) = 0;
    virtual cudaError_t Push(int level, unsigned long long queue_len, unsigned long long edges_hint) = 0;
    virtual cudaError_t Absorb(int level, int cur_out) = 0;
    virtual cudaError_t Pull(int level, int cur) = 0;
    virtual cudaError_t MergeVisited(int cur) = 0;
    virtual cudaError_t QueueToMap(int cur, unsigned long long queue_len) = 0;
    virtual cudaError_t MapToQueue(int cur) = 0;
    virtual cudaError_t Labels(int *labels_host) = 0;
    virtual cudaError_t ReadCounters(grb_mg_level *out) = 0;
    virtual int Launches() = 0;
};

template <typename SizeT>
struct MgRunner : MgRunnerBase {
    typedef gunrock::app::bfs::BFSProblem<int, SizeT, int, false, false> Problem;
    typedef gunrock::app::bfs::MGBFSEnactor<Problem> Enactor;
    Problem problem;
    Enactor enactor;
    cudaStream_t stream = 0;
    gunrock::util::PinnedArray<gunrock::oprtr::LevelCounters> h_counters;

    cudaError_t Create(gunrock::graph::DeviceCsr<int, SizeT> *g, cudaStream_t s)
    {
        stream = s;
        if (!g->HasInverse())
            return gunrock::util::GRError(cudaErrorInvalidValue, "partitioned BFS needs a symmetric graph", __FILE__, __LINE__);
        if (g->local_nodes % 32 != 0 && g->part.log_parts != 0)
            return gunrock::util::GRError(cudaErrorInvalidValue, "partition segments must be word aligned", __FILE__, __LINE__);
        GR_TRY(problem.Init(g, s));
        GR_TRY(problem.Reset(-1, gunrock::app::VERTEX_FRONTIERS, 1.0));    // sizes the queues
        GR_TRY(enactor.Setup(&problem, s));
        return h_counters.Allocate(1);
    }
    cudaError_t Reset(int src, unsigned long long *src_degree) override
    {
        GR_TRY(problem.Reset(src, gunrock::app::VERTEX_FRONTIERS, 1.0));
        GR_TRY(enactor.Reset(&problem));
        GR_TRY(cudaStreamSynchronize(stream));
        if (src_degree) *src_degree = problem.src_degree;
        return cudaSuccess;
    }
    cudaError_t Buffer(int which, void **ptr, unsigned long long *bytes) override
    {
        const unsigned long long seg_words = (unsigned long long)(problem.graph->local_nodes >> 5);
        const unsigned long long map_bytes = 4ull * seg_words * (unsigned long long)problem.graph->part.Parts();
        switch (which) {
            case 0: *ptr = problem.visited_mask.ptr;     *bytes = map_bytes; break;
            case 1: *ptr = problem.frontier_map[0].ptr;  *bytes = map_bytes; break;
            case 2: *ptr = problem.frontier_map[1].ptr;  *bytes = map_bytes; break;
            case 3: *ptr = problem.next_map.ptr;         *bytes = map_bytes; break;
            case 4: *ptr = problem.recv_map.ptr;         *bytes = map_bytes; break;
            case 5: *ptr = problem.labels.ptr;           *bytes = 4ull * (unsigned long long)problem.graph->local_nodes; break;
            case 6: *ptr = enactor.Counters();           *bytes = sizeof(gunrock::oprtr::LevelCounters); break;
            default: return cudaErrorInvalidValue;
        }
        return cudaSuccess;
    }
    cudaError_t Push(int level, unsigned long long queue_len, unsigned long long edges_hint) override
    { return enactor.Push(&problem, level, queue_len, edges_hint); }
    cudaError_t Absorb(int level, int cur_out) override { return enactor.Absorb(&problem, level, cur_out & 1); }
    cudaError_t Pull(int level, int cur) override { return enactor.Pull(&problem, level, cur & 1); }
    cudaError_t MergeVisited(int cur) override { return enactor.MergeVisited(&problem, cur & 1); }
    cudaError_t QueueToMap(int cur, unsigned long long queue_len) override { return enactor.QueueToMap(&problem, cur & 1, queue_len); }
    cudaError_t MapToQueue(int cur) override { return enactor.MapToQueue(&problem, cur & 1); }
    cudaError_t Labels(int *labels_host) override { return problem.Extract(labels_host, nullptr); }
    cudaError_t ReadCounters(grb_mg_level *out) override
    {
        GR_TRY(cudaMemcpyAsync(h_counters.ptr, enactor.Counters(), sizeof(gunrock::oprtr::LevelCounters),
                               cudaMemcpyDeviceToHost, stream));
        GR_TRY(cudaStreamSynchronize(stream));
        out->frontier_local = h_counters.ptr->out_len;  out->frontier_edges = h_counters.ptr->out_degree;
        out->edges_inspected = h_counters.ptr->edges;   out->candidates = h_counters.ptr->aux0;
        out->overflow = h_counters.ptr->overflow;
        return cudaSuccess;
    }
    int Launches() override { return enactor.enactor_stats.launches; }
};

struct grb_mgbfs_s {
    grb_graph_t graph = nullptr;
    MgRunnerBase *runner = nullptr;
    ~grb_mgbfs_s() { delete runner; }
};

#define MG_ENTER(ctx)                                         \
    if (!(ctx)) return (int)cudaErrorInvalidValue;            \
    GRB_TRY(cudaSetDevice((ctx)->graph->device()))

extern "C" int grb_mgbfs_create(grb_graph_t g, grb_mgbfs_t *out)
{
    if (!g || !out) return (int)cudaErrorInvalidValue;
    *out = nullptr;
    GRB_TRY(cudaSetDevice(g->device()));
    grb_mgbfs_s *c = new grb_mgbfs_s();
    c->graph = g;
    cudaError_t e;
    if (g->g32) { auto *r = new MgRunner<int>();       c->runner = r; e = r->Create(g->g32, g->stream); }
    else        { auto *r = new MgRunner<long long>(); c->runner = r; e = r->Create(g->g64, g->stream); }
    if (e != cudaSuccess) { delete c; return (int)e; }
    *out = c;
    return 0;
}
extern "C" int grb_mgbfs_reset(grb_mgbfs_t ctx, int32_t src, uint64_t *src_degree)
{
    MG_ENTER(ctx);
    unsigned long long d = 0;
    GRB_TRY(ctx->runner->Reset(src, &d));
    if (src_degree) *src_degree = d;
    return 0;
}
extern "C" int grb_mgbfs_buffer(grb_mgbfs_t ctx, int which, void **dev_ptr, uint64_t *bytes)
{
    if (!ctx || !dev_ptr || !bytes) return (int)cudaErrorInvalidValue;
    unsigned long long b = 0;
    GRB_TRY(ctx->runner->Buffer(which, dev_ptr, &b));
    *bytes = b;
    return 0;
}
extern "C" int grb_mgbfs_push(grb_mgbfs_t ctx, int32_t level, uint64_t queue_len, uint64_t edges_hint)
{ MG_ENTER(ctx); GRB_TRY(ctx->runner->Push(level, queue_len, edges_hint)); return 0; }
extern "C" int grb_mgbfs_absorb(grb_mgbfs_t ctx, int32_t level, int cur_out)
{ MG_ENTER(ctx); GRB_TRY(ctx->runner->Absorb(level, cur_out)); return 0; }
extern "C" int grb_mgbfs_pull(grb_mgbfs_t ctx, int32_t level, int cur)
{ MG_ENTER(ctx); GRB_TRY(ctx->runner->Pull(level, cur)); return 0; }
extern "C" int grb_mgbfs_merge_visited(grb_mgbfs_t ctx, int cur)
{ MG_ENTER(ctx); GRB_TRY(ctx->runner->MergeVisited(cur)); return 0; }
extern "C" int grb_mgbfs_queue_to_map(grb_mgbfs_t ctx, int cur, uint64_t queue_len)
{ MG_ENTER(ctx); GRB_TRY(ctx->runner->QueueToMap(cur, queue_len)); return 0; }
extern "C" int grb_mgbfs_map_to_queue(grb_mgbfs_t ctx, int cur)
{ MG_ENTER(ctx); GRB_TRY(ctx->runner->MapToQueue(cur)); return 0; }
extern "C" int grb_mgbfs_counters(grb_mgbfs_t ctx, grb_mg_level *out)
{
    MG_ENTER(ctx);
    if (!out) return (int)cudaErrorInvalidValue;
    memset(out, 0, sizeof(*out));
    GRB_TRY(ctx->runner->ReadCounters(out));
    out->gpu_launches = (uint64_t)ctx->runner->Launches();
    return 0;
}
extern "C" int grb_mgbfs_sync(grb_mgbfs_t ctx)
{ MG_ENTER(ctx); GRB_TRY(cudaStreamSynchronize(ctx->graph->stream)); return 0; }
extern "C" int grb_mgbfs_stream(grb_mgbfs_t ctx, void **cuda_stream)
{
    if (!ctx || !cuda_stream) return (int)cudaErrorInvalidValue;
    *cuda_stream = (void *)ctx->graph->stream;
    return 0;
}
extern "C" int grb_mgbfs_labels(grb_mgbfs_t ctx, int32_t *labels_host)
{ MG_ENTER(ctx); GRB_TRY(ctx->runner->Labels(labels_host)); return 0; }
extern "C" void grb_mgbfs_destroy(grb_mgbfs_t ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->graph->device());
    delete ctx;
}
