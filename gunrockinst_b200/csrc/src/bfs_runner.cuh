// bfs_runner.cuh -- glue between the C ABI and the BFS Problem / Enactor templates.
// Counterpart of run_bfs<VertexId, Value, SizeT, MARK_PREDECESSORS, ENABLE_IDEMPOTENCE>
// (gunrock/app/bfs/bfs_app.cu:146-229): Init -> Reset -> Enact -> Extract, with the
// Init done once per context instead of once per call.
#pragma once
#include "../../../include/gunrock_b200.h"
#include "graph_handle.cuh"
#include "../gunrock/app/bfs/bfs_enactor.cuh"

struct BfsRunnerBase {
    virtual ~BfsRunnerBase() {}
    virtual cudaError_t Run(int src, int *labels_host, int *preds_host, grb_stats *stats) = 0;
    virtual cudaError_t RunMany(const int *srcs, int count, grb_stats *stats) = 0;
    virtual void *DeviceLabels() = 0;
    virtual const std::vector<gunrock::app::bfs::LevelRecord> &Levels() = 0;
};

template <typename SizeT, bool MARK_PRED, bool IDEM>
struct BfsRunner : BfsRunnerBase {
    typedef gunrock::app::bfs::BFSProblem<int, SizeT, int, MARK_PRED, IDEM> Problem;
    typedef gunrock::app::bfs::BFSEnactor<Problem> Enactor;
    Problem problem;
    Enactor enactor;
    gunrock::app::bfs::BFSOptions opt;
    double queue_sizing = 1.0;
    cudaStream_t stream = 0;
    cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr;
    bool ready = false;
    // degree order: the problem runs on the graph's twin; sources are translated on the way in,
    // labels / predecessors on the way out (lazily: only when somebody asks for them)
    const int *to_new = nullptr, *to_old = nullptr;

    ~BfsRunner()
    {
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
        if (e2) cudaEventDestroy(e2);
    }

    cudaError_t Create(gunrock::graph::DeviceCsr<int, SizeT> *g, const grb_bfs_config *cfg, cudaStream_t s,
                       const int *map_to_new = nullptr, const int *map_to_old = nullptr)
    {
        stream = s;
        to_new = map_to_new; to_old = map_to_old;
        if (!cfg && !map_to_new) opt.stream_min_edges = 0;
        if (cfg) {
            opt.traversal_mode = cfg->traversal_mode;
            opt.direction_optimized = cfg->direction_optimized != 0;
            if (cfg->alpha > 0) opt.alpha = cfg->alpha;
            if (cfg->beta > 0) opt.beta = cfg->beta;
            if (cfg->queue_sizing > 0) queue_sizing = cfg->queue_sizing;
            opt.record_levels = cfg->record_levels != 0;
            // 0: choose by size -- one persistent kernel below 2^24 vertices (a level costs less than a
            // launch + counter fetch there), the host-driven loop of stand-alone kernels above (at scale
            // 25 they run the big levels 1.1-1.35x faster: 5 instead of 4 resident CTAs per SM, fewer
            // live registers around the operator bodies); 1: host loop; 2: persistent kernel
            // streaming advance / filter: 0 = by default on degree-ordered twins only, 1 = off, n > 1 = from n entries upward
            if (cfg->stream_min_edges == 1 || (cfg->stream_min_edges == 0 && !map_to_new)) opt.stream_min_edges = 0;
            else if (cfg->stream_min_edges > 1) opt.stream_min_edges = (unsigned long long)cfg->stream_min_edges;
            opt.host_loop = cfg->host_loop == 1 || (cfg->host_loop == 0 && g->nodes >= (1ll << 24));
        }
        GR_TRY(problem.Init(g, s));
        GR_TRY(problem.SetTwinMaps(to_new, to_old));
        GR_TRY(cudaEventCreate(&e0)); GR_TRY(cudaEventCreate(&e1)); GR_TRY(cudaEventCreate(&e2));
        return cudaSuccess;
    }

    cudaError_t Run(int src, int *labels_host, int *preds_host, grb_stats *stats) override
    {
        using gunrock::app::VERTEX_FRONTIERS;
        using gunrock::app::EDGE_FRONTIERS;
        if (to_new && src >= 0 && (long long)src < problem.graph->nodes) {      // the twin's id of the source
            GR_TRY(cudaMemcpyAsync(&src, to_new + src, sizeof(int), cudaMemcpyDeviceToHost, stream));
            GR_TRY(cudaStreamSynchronize(stream));
        }
        GR_TRY(cudaEventRecord(e0, stream));
        const bool fused_reset = ready && enactor.FusesReset(src, &problem, opt);
        GR_TRY(problem.Reset(src, IDEM ? EDGE_FRONTIERS : VERTEX_FRONTIERS, queue_sizing, fused_reset));
        if (!ready) { GR_TRY(enactor.Setup(&problem, stream)); ready = true; }
        GR_TRY(cudaEventRecord(e1, stream));
        GR_TRY(enactor.Enact(&problem, src, opt, fused_reset));
        GR_TRY(cudaEventRecord(e2, stream));
        GR_TRY(cudaEventSynchronize(e2));
        if (IDEM) GR_TRY(enactor.RecountEdgesVisited(&problem));
        if (labels_host) GR_TRY(Extract(labels_host, preds_host));
        if (stats) {
            float r = 0, t = 0;
            cudaEventElapsedTime(&r, e0, e1); cudaEventElapsedTime(&t, e1, e2);
            memset(stats, 0, sizeof(*stats));
            stats->reset_ms = r; stats->elapsed_ms = t;
            stats->iterations = enactor.enactor_stats.iteration;
            stats->edges_visited = (int64_t)enactor.edges_visited;
            stats->edges_inspected = enactor.enactor_stats.edges_inspected;
            stats->total_queued = enactor.enactor_stats.total_queued;
            stats->nodes_visited = (int64_t)enactor.enactor_stats.total_queued;
            stats->algo_bytes = enactor.enactor_stats.algo_bytes;
            stats->push_levels = enactor.enactor_stats.push_levels;
            stats->pull_levels = enactor.enactor_stats.pull_levels;
            stats->gpu_launches = enactor.enactor_stats.launches;
            stats->reserved = enactor.kernel_us;        // persistent loop: microseconds between the kernel's first and last instruction
        }
        return cudaSuccess;
    }
    // reference: BFSProblem::Extract (bfs_problem.cuh:144-177); on a twin the search itself ended with the
    // gather into the caller's ids (bfs_enactor.cuh / bfs_persistent.cuh)
    cudaError_t Extract(int *labels_host, int *preds_host)
    {
        if (!to_new) return problem.Extract(labels_host, preds_host);
        const size_t n = (size_t)problem.graph->nodes;
        GR_TRY(problem.out_labels.ToHost(labels_host, n, stream));
        if (MARK_PRED && preds_host) GR_TRY(problem.out_preds.ToHost(preds_host, n, stream));
        return cudaStreamSynchronize(stream);
    }
    // `count` searches; when the persistent kernel can take them, batches of up to kBatch sources run
    // as ONE launch each (no launch, copy or host synchronisation between two searches)
    cudaError_t RunMany(const int *srcs, int count, grb_stats *stats) override
    {
        using gunrock::app::VERTEX_FRONTIERS;
        const int kBatch = 256;
        int done = 0;
        while (done < count) {
            bool batch = ready && !IDEM && !opt.record_levels && count - done > 1;
            int take = count - done < kBatch ? count - done : kBatch;
            for (int i = 0; batch && i < take; ++i) batch = enactor.FusesReset(srcs[done + i], &problem, opt);
                if (!batch) {
                GR_TRY(Run(srcs[done], nullptr, nullptr, stats ? stats + done : nullptr));
                ++done;
                continue;
            }
            if constexpr (!IDEM) {
                std::vector<gunrock::app::bfs::PersistentSummary> sums((size_t)take);
                GR_TRY(problem.Reset(srcs[done + take - 1], VERTEX_FRONTIERS, queue_sizing, true));
                cudaError_t be = enactor.Enact(&problem, srcs[done + take - 1], opt, true, srcs + done, take, sums.data());
                if (be == cudaErrorCooperativeLaunchTooLarge && !enactor.persistent_ok) continue;   // search by search from here on
                GR_TRY(be);
                for (int i = 0; stats && i < take; ++i) {
                    grb_stats *st = stats + done + i;
                    memset(st, 0, sizeof(*st));
                    st->elapsed_ms = (float)((double)(sums[i].ns_kernel - sums[i].ns_reset) * 1e-6);   // device-side spans
                    st->reset_ms = (float)((double)sums[i].ns_reset * 1e-6);
                    st->iterations = (int64_t)sums[i].levels;
                    st->edges_visited = (int64_t)sums[i].edges_visited;
                    st->edges_inspected = sums[i].edges_inspected;
                    st->total_queued = sums[i].total_queued;
                    st->nodes_visited = (int64_t)sums[i].total_queued;
                    st->algo_bytes = sums[i].algo_bytes;
                    st->push_levels = (int32_t)sums[i].push_levels;
                    st->pull_levels = (int32_t)sums[i].pull_levels;
                    st->gpu_launches = i == 0 ? 1 : 0;             // one kernel for the whole batch
                    st->reserved = (int32_t)(sums[i].ns_kernel / 1000);
                }
            }
            done += take;
        }
        return cudaSuccess;
    }
    void *DeviceLabels() override { return to_new ? problem.out_labels.ptr : problem.labels.ptr; }
    const std::vector<gunrock::app::bfs::LevelRecord> &Levels() override { return enactor.levels; }
};

// to_new / to_old non-NULL: `g` is the degree-ordered twin of the caller's graph
BfsRunnerBase *MakeBfsRunner32(Csr32 *g, const grb_bfs_config *cfg, cudaStream_t s, cudaError_t *err,
                               const int *to_new = nullptr, const int *to_old = nullptr);
BfsRunnerBase *MakeBfsRunner64(Csr64 *g, const grb_bfs_config *cfg, cudaStream_t s, cudaError_t *err,
                               const int *to_new = nullptr, const int *to_old = nullptr);

template <typename R, typename G>
static BfsRunnerBase *MakeOne(G *g, const grb_bfs_config *cfg, cudaStream_t s, cudaError_t *err,
                              const int *to_new, const int *to_old)
{
    R *r = new R();
    *err = r->Create(g, cfg, s, to_new, to_old);
    if (*err != cudaSuccess) { delete r; return nullptr; }
    return r;
}
