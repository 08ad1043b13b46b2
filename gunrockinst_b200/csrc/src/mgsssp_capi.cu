// mgsssp_capi.cu -- C entry points of the 1D-partitioned SSSP (include/gunrock_b200.h,
// grb_mgsssp_*).  No reference counterpart: gunrock/app/problem_base.cuh:336-338 leaves the
// multi-GPU path as a TODO; relaxation arithmetic and output conventions are the reference's
// (gunrock/app/sssp/sssp_functor.cuh:52-66, sssp_problem.cuh:325,372).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/app/sssp/mgsssp_enactor.cuh"

struct MgSsspBase {
    virtual ~MgSsspBase() {}
    virtual cudaError_t PeerExport(unsigned char *handles) = 0;
    virtual cudaError_t PeerImport(int world, const unsigned char *all) = 0;
    virtual cudaError_t Run(int src, grb_mg_stats *out) = 0;
    virtual cudaError_t Result(unsigned *dist, int *preds) = 0;
};

template <typename SizeT>
struct MgSsspRunner : MgSsspBase {
    gunrock::app::sssp::MGSSSPEnactor<SizeT> enactor;
    std::vector<void *> opened;

    void *PeerBuffer(int which)
    {
        switch (which) {
            case 0: return enactor.dist_pred.ptr;
            case 1: return enactor.touched.ptr;
            default: return enactor.comm.mailbox.ptr;
        }
    }
    cudaError_t PeerExport(unsigned char *handles) override
    {
        for (int b = 0; b < GRB_MGSSSP_PEER_BUFFERS; ++b) {
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
            void *ptrs[GRB_MGSSSP_PEER_BUFFERS];
            for (int b = 0; b < GRB_MGSSSP_PEER_BUFFERS; ++b) {
                cudaIpcMemHandle_t h;
                memcpy(&h, all + ((size_t)r * GRB_MGSSSP_PEER_BUFFERS + b) * GRB_IPC_HANDLE_BYTES, sizeof(h));
                GR_TRY(cudaIpcOpenMemHandle(&ptrs[b], h, cudaIpcMemLazyEnablePeerAccess));
                opened.push_back(ptrs[b]);
            }
            enactor.peers.dist_pred[r] = (unsigned long long *)ptrs[0];
            enactor.peers.touched[r]   = (unsigned *)ptrs[1];
            mt.mailbox[r] = (gunrock::app::bfs::Mailbox *)ptrs[2];
        }
        enactor.comm.ready = true;
        return cudaSuccess;
    }
    cudaError_t Run(int src, grb_mg_stats *out) override
    {
        typename gunrock::app::sssp::MGSSSPEnactor<SizeT>::SearchStats st;
        GR_TRY(enactor.Run(src, &st));
        if (out) {
            memset(out, 0, sizeof(*out));
            out->elapsed_ms = st.elapsed_ms; out->iterations = st.iterations; out->gpu_launches = st.launches;
            out->edges_inspected = st.relaxations; out->nodes_visited = st.queued;
        }
        return cudaSuccess;
    }
    cudaError_t Result(unsigned *dist, int *preds) override { return enactor.Extract(dist, preds); }
    ~MgSsspRunner() override { for (void *p : opened) cudaIpcCloseMemHandle(p); }
};

struct grb_mgsssp_s {
    grb_graph_t graph = nullptr;
    MgSsspBase *runner = nullptr;
    ~grb_mgsssp_s() { delete runner; }
};

#define MGS_ENTER(ctx)                                        \
    if (!(ctx)) return (int)cudaErrorInvalidValue;            \
    GRB_TRY(cudaSetDevice((ctx)->graph->device()))

extern "C" int grb_mgsssp_create(grb_graph_t g, grb_mgsssp_t *out)
{
    if (!g || !out) return (int)cudaErrorInvalidValue;
    *out = nullptr;
    GRB_TRY(cudaSetDevice(g->device()));
    grb_mgsssp_s *c = new grb_mgsssp_s();
    c->graph = g;
    cudaError_t e;
    if (g->g32) { auto *r = new MgSsspRunner<int>();       c->runner = r; e = r->enactor.Setup(g->g32, g->stream); }
    else        { auto *r = new MgSsspRunner<long long>(); c->runner = r; e = r->enactor.Setup(g->g64, g->stream); }
    if (e != cudaSuccess) { delete c; return (int)e; }
    *out = c;
    return 0;
}
extern "C" int grb_mgsssp_peer_export(grb_mgsssp_t ctx, unsigned char *handles)
{ MGS_ENTER(ctx); if (!handles) return (int)cudaErrorInvalidValue; GRB_TRY(ctx->runner->PeerExport(handles)); return 0; }
extern "C" int grb_mgsssp_peer_import(grb_mgsssp_t ctx, int world, const unsigned char *all_handles)
{
    MGS_ENTER(ctx);
    if (!all_handles || world < 1) return (int)cudaErrorInvalidValue;
    GRB_TRY(ctx->runner->PeerImport(world, all_handles));
    return 0;
}
extern "C" int grb_mgsssp_run(grb_mgsssp_t ctx, int32_t src, grb_mg_stats *out)
{ MGS_ENTER(ctx); GRB_TRY(ctx->runner->Run(src, out)); return 0; }
extern "C" int grb_mgsssp_result(grb_mgsssp_t ctx, uint32_t *dist_host, int32_t *preds_host)
{ MGS_ENTER(ctx); GRB_TRY(ctx->runner->Result(dist_host, preds_host)); return 0; }
extern "C" void grb_mgsssp_destroy(grb_mgsssp_t ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->graph->device());
    delete ctx;
}
