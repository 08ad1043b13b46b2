// mgbfs_capi.cu -- C entry points of the 1D-partitioned BFS steps (include/gunrock_b200.h,
// grb_mgbfs_*).  No reference counterpart: gunrock/app/problem_base.cuh:336-338 leaves the
// multi-GPU path as a TODO; the striping rule it declares (:186-210) is the one used here.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
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
    virtual cudaError_t PeerExport(unsigned char *handles) = 0;
    virtual cudaError_t PeerImport(int world, const unsigned char *all, unsigned long long global_edges) = 0;
    virtual cudaError_t RunPeer(int src, bool dopt, double alpha, double beta, bool record, grb_mg_stats *out) = 0;
    virtual int PeerLevels(grb_mg_trace *out, int cap) = 0;
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
        // a directed slice without a transpose is searched top-down only (no pull levels): see RunPeer
        if (g->local_nodes % 32 != 0 && g->part.log_parts != 0)
            return gunrock::util::GRError(cudaErrorInvalidValue, "partition segments must be word aligned", __FILE__, __LINE__);
        GR_TRY(problem.Init(g, s));
        GR_TRY(problem.Reset(-1, gunrock::app::VERTEX_FRONTIERS, 1.0));    // sizes the queues
        GR_TRY(enactor.Setup(&problem, s));
        GR_TRY(enactor.SetupPeers(&problem));
        enactor.global_edges = (unsigned long long)g->edges;     // one device; peer import sets the global sum
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

    // ---- peer-memory exchange ----
    std::vector<void *> opened;         // IPC mappings to close
    void *PeerBuffer(int which)
    {
        switch (which) {
            case 0: return problem.frontier_map[0].ptr;
            case 1: return problem.frontier_map[1].ptr;
            case 2: return problem.recv_map.ptr;
            default: return enactor.mailbox.ptr;
        }
    }
    cudaError_t PeerExport(unsigned char *handles) override
    {
        static_assert(sizeof(cudaIpcMemHandle_t) == GRB_IPC_HANDLE_BYTES, "IPC handle size");
        for (int b = 0; b < GRB_MG_PEER_BUFFERS; ++b) {
            cudaIpcMemHandle_t h;
            GR_TRY(cudaIpcGetMemHandle(&h, PeerBuffer(b)));
            memcpy(handles + b * GRB_IPC_HANDLE_BYTES, &h, sizeof(h));
        }
        return cudaSuccess;
    }
    cudaError_t PeerImport(int world, const unsigned char *all, unsigned long long global_edges) override
    {
        gunrock::app::bfs::PeerTable &pt = enactor.peers;
        if (world != pt.parts)
            return gunrock::util::GRError(cudaErrorInvalidValue, "peer import: world size differs from the partition", __FILE__, __LINE__);
        enactor.global_edges = global_edges;
        for (int r = 0; r < world; ++r) {
            if (r == pt.rank) continue;
            void *ptrs[GRB_MG_PEER_BUFFERS];
            for (int b = 0; b < GRB_MG_PEER_BUFFERS; ++b) {
                cudaIpcMemHandle_t h;
                memcpy(&h, all + ((size_t)r * GRB_MG_PEER_BUFFERS + b) * GRB_IPC_HANDLE_BYTES, sizeof(h));
                GR_TRY(cudaIpcOpenMemHandle(&ptrs[b], h, cudaIpcMemLazyEnablePeerAccess));
                opened.push_back(ptrs[b]);
            }
            pt.map[0][r] = (unsigned *)ptrs[0];
            pt.map[1][r] = (unsigned *)ptrs[1];
            pt.recv[r]   = (unsigned *)ptrs[2];
            pt.mailbox[r] = (gunrock::app::bfs::Mailbox *)ptrs[3];
        }
        enactor.p2p_ready = true;
        return cudaSuccess;
    }
    cudaError_t RunPeer(int src, bool dopt, double alpha, double beta, bool record, grb_mg_stats *out) override
    {
        typename Enactor::SearchStats st;
        GR_TRY(enactor.RunPeer(&problem, src, dopt, alpha, beta, record, &st));
        if (out) {
            out->elapsed_ms = st.elapsed_ms; out->iterations = st.iterations; out->push_levels = st.push_levels;
            out->pull_levels = st.pull_levels; out->gpu_launches = st.launches; out->edges_visited = st.edges_visited;
            out->edges_inspected = st.edges_inspected; out->exchange_bytes = st.exchange_bytes;
            out->nodes_visited = st.nodes_visited;
        }
        return cudaSuccess;
    }
    int PeerLevels(grb_mg_trace *out, int cap) override
    {
        int k = 0;
        for (const auto &t : enactor.trace) {
            if (k >= cap) break;
            out[k].level = t.level; out[k].kind = t.kind; out[k].frontier = t.frontier; out[k].discovered = t.discovered;
            out[k].edges = t.edges; out[k].max_edges = t.max_edges; out[k].exchange_bytes = t.exchange_bytes; out[k].candidates = t.candidates;
            out[k].compute_ms = t.compute_ms; out[k].exchange_ms = t.exchange_ms;
            ++k;
        }
        return k;
    }
    ~MgRunner() override
    {
        for (void *p : opened) cudaIpcCloseMemHandle(p);
        enactor.DestroyEvents();
    }
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
extern "C" int grb_mgbfs_peer_export(grb_mgbfs_t ctx, unsigned char *handles)
{ MG_ENTER(ctx); if (!handles) return (int)cudaErrorInvalidValue; GRB_TRY(ctx->runner->PeerExport(handles)); return 0; }
extern "C" int grb_mgbfs_peer_import(grb_mgbfs_t ctx, int world, const unsigned char *all_handles, uint64_t global_edges)
{
    MG_ENTER(ctx);
    if (!all_handles || world < 1) return (int)cudaErrorInvalidValue;
    GRB_TRY(ctx->runner->PeerImport(world, all_handles, global_edges));
    return 0;
}
extern "C" int grb_mgbfs_run_peer(grb_mgbfs_t ctx, int32_t src, int direction_optimized, double alpha, double beta,
                                  int record_levels, grb_mg_stats *out)
{
    MG_ENTER(ctx);
    if (out) memset(out, 0, sizeof(*out));
    GRB_TRY(ctx->runner->RunPeer(src, direction_optimized != 0, alpha, beta, record_levels != 0, out));
    return 0;
}
extern "C" int grb_mgbfs_peer_levels(grb_mgbfs_t ctx, grb_mg_trace *out, int32_t cap, int32_t *count)
{
    if (!ctx || !out || !count) return (int)cudaErrorInvalidValue;
    *count = ctx->runner->PeerLevels(out, cap);
    return 0;
}
extern "C" void grb_mgbfs_destroy(grb_mgbfs_t ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->graph->device());
    delete ctx;
}
