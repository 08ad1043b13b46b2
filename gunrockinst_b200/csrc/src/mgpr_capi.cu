// mgpr_capi.cu -- C entry points of the 1D-partitioned PageRank (include/gunrock_b200.h,
// grb_mgpr_*).  No reference counterpart: gunrock/app/problem_base.cuh:336-338 leaves the
// multi-GPU path as a TODO; the arithmetic and the pruning are the reference's
// (gunrock/app/pr/pr_functor.cuh:32-175, pr_enactor.cuh:247-329,346-487).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/app/pr/mgpr_enactor.cuh"

struct MgPrBase {
    virtual ~MgPrBase() {}
    virtual cudaError_t PeerExport(unsigned char *handles) = 0;
    virtual cudaError_t PeerImport(int world, const unsigned char *all) = 0;
    virtual cudaError_t Run(int src, double delta, double threshold, long long max_iter, grb_mg_stats *out) = 0;
    virtual cudaError_t Result(double *ranks) = 0;
};

template <typename SizeT>
struct MgPrRunner : MgPrBase {
    gunrock::app::pr::MGPREnactor<SizeT> enactor;
    std::vector<void *> opened;

    void *PeerBuffer(int which)
    {
        switch (which) {
            case 0: return enactor.contrib.ptr;
            case 1: return enactor.zero.ptr;
            default: return enactor.comm.mailbox.ptr;
        }
    }
    cudaError_t PeerExport(unsigned char *handles) override
    {
        for (int b = 0; b < GRB_MGPR_PEER_BUFFERS; ++b) {
            cudaIpcMemHandle_t h;
            GR_TRY(cudaIpcGetMemHandle(&h, PeerBuffer(b)));
            memcpy(handles + b * GRB_IPC_HANDLE_BYTES, &h, sizeof(h));
        }
        return cudaSuccess;
    }
    cudaError_t PeerImport(int world, const unsigned char *all) override
    {
        auto &mt = enactor.comm.table;
        if (world != mt.parts)
            return gunrock::util::GRError(cudaErrorInvalidValue, "peer import: world size differs from the partition", __FILE__, __LINE__);
        for (int r = 0; r < world; ++r) {
            if (r == mt.rank) continue;
            void *ptrs[GRB_MGPR_PEER_BUFFERS];
            for (int b = 0; b < GRB_MGPR_PEER_BUFFERS; ++b) {
                cudaIpcMemHandle_t h;
                memcpy(&h, all + ((size_t)r * GRB_MGPR_PEER_BUFFERS + b) * GRB_IPC_HANDLE_BYTES, sizeof(h));
                GR_TRY(cudaIpcOpenMemHandle(&ptrs[b], h, cudaIpcMemLazyEnablePeerAccess));
                opened.push_back(ptrs[b]);
            }
            enactor.peers.contrib[r] = (double *)ptrs[0];
            enactor.peers.zero[r]    = (unsigned *)ptrs[1];
            mt.mailbox[r] = (gunrock::app::bfs::Mailbox *)ptrs[2];
        }
        enactor.comm.ready = true;
        return cudaSuccess;
    }
    cudaError_t Run(int src, double delta, double threshold, long long max_iter, grb_mg_stats *out) override
    {
        typename gunrock::app::pr::MGPREnactor<SizeT>::RunStats st;
        GR_TRY(enactor.Run(src, delta, threshold, max_iter, &st));
        if (out) {
            memset(out, 0, sizeof(*out));
            out->elapsed_ms = st.elapsed_ms; out->iterations = st.iterations; out->gpu_launches = st.launches;
            out->push_levels = st.prune_sweeps;
            out->edges_inspected = st.edges; out->nodes_visited = st.valid_nodes; out->exchange_bytes = st.exchange_bytes;
            out->edges_visited = (uint64_t)(st.sweep_ms * 1e3);
        }
        return cudaSuccess;
    }
    cudaError_t Result(double *ranks) override { return enactor.Extract(ranks); }
    ~MgPrRunner() override { for (void *p : opened) cudaIpcCloseMemHandle(p); }
};

struct grb_mgpr_s {
    grb_graph_t graph = nullptr;
    MgPrBase *runner = nullptr;
    ~grb_mgpr_s() { delete runner; }
};

#define MGP_ENTER(ctx)                                        \
    if (!(ctx)) return (int)cudaErrorInvalidValue;            \
    GRB_TRY(cudaSetDevice((ctx)->graph->device()))

extern "C" int grb_mgpr_create(grb_graph_t g, grb_mgpr_t *out)
{
    if (!g || !out) return (int)cudaErrorInvalidValue;
    *out = nullptr;
    GRB_TRY(cudaSetDevice(g->device()));
    grb_mgpr_s *c = new grb_mgpr_s();
    c->graph = g;
    cudaError_t e;
    if (g->g32) { auto *r = new MgPrRunner<int>();       c->runner = r; e = r->enactor.Setup(g->g32, g->stream); }
    else        { auto *r = new MgPrRunner<long long>(); c->runner = r; e = r->enactor.Setup(g->g64, g->stream); }
    if (e != cudaSuccess) { delete c; return (int)e; }
    *out = c;
    return 0;
}
extern "C" int grb_mgpr_peer_export(grb_mgpr_t ctx, unsigned char *handles)
{ MGP_ENTER(ctx); if (!handles) return (int)cudaErrorInvalidValue; GRB_TRY(ctx->runner->PeerExport(handles)); return 0; }
extern "C" int grb_mgpr_peer_import(grb_mgpr_t ctx, int world, const unsigned char *all_handles)
{
    MGP_ENTER(ctx);
    if (!all_handles || world < 1) return (int)cudaErrorInvalidValue;
    GRB_TRY(ctx->runner->PeerImport(world, all_handles));
    return 0;
}
extern "C" int grb_mgpr_run(grb_mgpr_t ctx, int32_t src, double delta, double threshold, int32_t max_iter, grb_mg_stats *out)
{ MGP_ENTER(ctx); GRB_TRY(ctx->runner->Run(src, delta, threshold, max_iter, out)); return 0; }
extern "C" int grb_mgpr_result(grb_mgpr_t ctx, double *ranks_host)
{ MGP_ENTER(ctx); if (!ranks_host) return (int)cudaErrorInvalidValue; GRB_TRY(ctx->runner->Result(ranks_host)); return 0; }
extern "C" void grb_mgpr_destroy(grb_mgpr_t ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->graph->device());
    delete ctx;
}
