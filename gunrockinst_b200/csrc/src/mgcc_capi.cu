// mgcc_capi.cu -- C entry points of the 1D-partitioned connected components (include/gunrock_b200.h,
// grb_mgcc_*).  No reference counterpart: gunrock/app/problem_base.cuh:336-338 leaves the
// multi-GPU path as a TODO; the output convention (label = smallest vertex id of the component)
// is the one the single-device grb_cc / gunrock_cc_func produce (gunrock/app/cc/cc_problem.cuh:164-170).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/app/cc/mgcc_enactor.cuh"

struct MgCcBase {
    virtual ~MgCcBase() {}
    virtual cudaError_t PeerExport(unsigned char *handles) = 0;
    virtual cudaError_t PeerImport(int world, const unsigned char *all) = 0;
    virtual cudaError_t Run(grb_mg_stats *out) = 0;
    virtual cudaError_t Result(int *labels, long long *roots) = 0;
};

template <typename SizeT>
struct MgCcRunner : MgCcBase {
    gunrock::app::cc::MGCCEnactor<SizeT> enactor;
    std::vector<void *> opened;

    void *PeerBuffer(int which)
    {
        switch (which) {
            case 0: return enactor.label_word.ptr;
            case 1: return enactor.touched.ptr;
            default: return enactor.comm.mailbox.ptr;
        }
    }
    cudaError_t PeerExport(unsigned char *handles) override
    {
        for (int b = 0; b < GRB_MGCC_PEER_BUFFERS; ++b) {
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
            void *ptrs[GRB_MGCC_PEER_BUFFERS];
            for (int b = 0; b < GRB_MGCC_PEER_BUFFERS; ++b) {
                cudaIpcMemHandle_t h;
                memcpy(&h, all + ((size_t)r * GRB_MGCC_PEER_BUFFERS + b) * GRB_IPC_HANDLE_BYTES, sizeof(h));
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
    cudaError_t Run(grb_mg_stats *out) override
    {
        typename gunrock::app::cc::MGCCEnactor<SizeT>::RunStats st;
        GR_TRY(enactor.Run(&st));
        if (out) {
            memset(out, 0, sizeof(*out));
            out->elapsed_ms = st.elapsed_ms; out->iterations = st.iterations; out->gpu_launches = st.launches;
            out->edges_inspected = st.updates; out->nodes_visited = st.queued;
        }
        return cudaSuccess;
    }
    cudaError_t Result(int *labels, long long *roots) override { return enactor.Extract(labels, roots); }
    ~MgCcRunner() override { for (void *p : opened) cudaIpcCloseMemHandle(p); }
};

struct grb_mgcc_s {
    grb_graph_t graph = nullptr;
    MgCcBase *runner = nullptr;
    ~grb_mgcc_s() { delete runner; }
};

#define MGC_ENTER(ctx)                                        \
    if (!(ctx)) return (int)cudaErrorInvalidValue;            \
    GRB_TRY(cudaSetDevice((ctx)->graph->device()))

extern "C" int grb_mgcc_create(grb_graph_t g, grb_mgcc_t *out)
{
    if (!g || !out) return (int)cudaErrorInvalidValue;
    *out = nullptr;
    GRB_TRY(cudaSetDevice(g->device()));
    grb_mgcc_s *c = new grb_mgcc_s();
    c->graph = g;
    cudaError_t e;
    if (g->g32) { auto *r = new MgCcRunner<int>();       c->runner = r; e = r->enactor.Setup(g->g32, g->stream); }
    else        { auto *r = new MgCcRunner<long long>(); c->runner = r; e = r->enactor.Setup(g->g64, g->stream); }
    if (e != cudaSuccess) { delete c; return (int)e; }
    *out = c;
    return 0;
}
extern "C" int grb_mgcc_peer_export(grb_mgcc_t ctx, unsigned char *handles)
{ MGC_ENTER(ctx); if (!handles) return (int)cudaErrorInvalidValue; GRB_TRY(ctx->runner->PeerExport(handles)); return 0; }
extern "C" int grb_mgcc_peer_import(grb_mgcc_t ctx, int world, const unsigned char *all_handles)
{
    MGC_ENTER(ctx);
    if (!all_handles || world < 1) return (int)cudaErrorInvalidValue;
    GRB_TRY(ctx->runner->PeerImport(world, all_handles));
    return 0;
}
extern "C" int grb_mgcc_run(grb_mgcc_t ctx, grb_mg_stats *out)
{ MGC_ENTER(ctx); GRB_TRY(ctx->runner->Run(out)); return 0; }
extern "C" int grb_mgcc_result(grb_mgcc_t ctx, int32_t *labels_host, int64_t *local_roots)
{
    MGC_ENTER(ctx);
    long long r = 0;
    GRB_TRY(ctx->runner->Result(labels_host, &r));
    if (local_roots) *local_roots = r;
    return 0;
}
extern "C" void grb_mgcc_destroy(grb_mgcc_t ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->graph->device());
    delete ctx;
}
