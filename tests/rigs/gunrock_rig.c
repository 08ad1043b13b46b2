/* gunrock_rig.c -- command-line test rig for the B200 build, shaped like the reference's
 * tests/{bfs,dobfs,sssp,cc,pr}/test_*.cu programs (usage text: tests/bfs/test_bfs.cu:58-89):
 *
 *   gunrock_rig <bfs|dobfs|sssp|cc|pr> market <file.mtx> [options]
 *   gunrock_rig <primitive> rmat --rmat_scale=<s> [--rmat_edgefactor=<f>] [--rmat_seed=<n>] [options]
 *
 *   --device=<n>  --undirected  --src=<id|randomize|largestdegree>  --quick[=<0|1>]  --v
 *   --iteration-num=<n>  --idempotence[=<0|1>]  --mark-pred  --queue-sizing=<f>
 *   --traversal-mode=<0|1|2>  --delta-factor=<n> (sssp)  --delta=<f> --error=<f> --max-iter=<n> (pr)
 *   --csr-cache-dir=<dir>   keep / reuse the reference's text cache of the built CSR (.<file>_undirected_csr /
 *                           _nonreversed_csr, gunrock/csr.cuh:140-232, graphio/market.cuh:254-262,312-326) in <dir>
 *                           (the reference writes it next to the .mtx; a scratch directory here)
 *
 * TEST INFRASTRUCTURE, not product: the graph is loaded by the oracle's restatement of the
 * reference reader (graphio/market.cuh:139-183, csr.cuh:247-340), the GPU side runs through the
 * C ABI of libgunrock.so (include/gunrock_b200.h), and -- unless --quick -- the result is checked
 * against the oracle's CPU algorithms the way the reference rigs check against theirs
 * (SimpleReferenceBfs tests/bfs/test_bfs.cu:258-322, Boost Dijkstra / connected_components /
 * page_rank call sites tests/{sssp,cc,pr}/test_*.cu).  Prints "... Validity: CORRECT" /
 * "INCORRECT" lines like gunrock/util/test_utils.cuh:311-337 and exits 0 only when every check
 * passed. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "gunrock_b200.h"

typedef int32_t vid_t;
typedef int64_t eid_t;
/* oracle/oracle.c */
int     orc_market_read(const char *path, int undirected, int reversed, int load_values,
                        int64_t *n_out, int64_t *ncoo_out, vid_t **rows, vid_t **cols, int64_t **vals);
int64_t orc_coo_to_csr(int64_t n, int64_t ncoo, const vid_t *rows, const vid_t *cols, const int64_t *vals,
                       eid_t *row_offsets, vid_t *col_out, int64_t *val_out);
int64_t orc_rmat_csr(int scale, int64_t edges, uint64_t seed, int undirected, double a, double b, double c, double d,
                     eid_t *row_offsets, vid_t *col_out);
void    orc_edge_weights(int64_t n, const eid_t *row_offsets, const vid_t *col, uint64_t seed, uint32_t *w);
int     orc_bfs(int64_t n, const eid_t *ro, const vid_t *col, vid_t src, vid_t *labels, vid_t *preds);
void    orc_sssp(int64_t n, const eid_t *ro, const vid_t *col, const uint32_t *w, vid_t src, uint32_t *dist, vid_t *preds);
int64_t orc_cc(int64_t n, const eid_t *ro, const vid_t *col, vid_t *labels);
int     orc_pagerank(int64_t n, const eid_t *ro, const vid_t *col, double delta, double threshold, int max_iter,
                     vid_t src, double *rank);
void    orc_free(void *p);
int     orc_csr_cache_write(const char *path, int64_t n, int64_t m, const eid_t *row_offsets, const vid_t *col, const int64_t *vals);
int     orc_csr_cache_read(const char *path, int load_values, int64_t *n_out, int64_t *m_out, eid_t **row_offsets, vid_t **col,
                           int64_t **vals);
double  orc_now_seconds(void);

static const char *opt_value(int argc, char **argv, const char *name)
{
    size_t k = strlen(name);
    for (int i = 1; i < argc; ++i) {
        if (strncmp(argv[i], "--", 2) == 0 && strncmp(argv[i] + 2, name, k) == 0) {
            if (argv[i][2 + k] == '=') return argv[i] + 3 + k;
            if (argv[i][2 + k] == 0) return "";
        }
    }
    return NULL;
}
static int opt_flag(int argc, char **argv, const char *name, int dflt)
{
    const char *v = opt_value(argc, argv, name);
    if (!v) return dflt;
    return v[0] == 0 ? 1 : atoi(v);
}

static void usage(void)
{
    printf(" gunrock_rig <bfs|dobfs|sssp|cc|pr> <graph type> <graph type args> [--device=<device_index>]\n"
           " (or, under the reference's names: test_bfs | test_dobfs | test_sssp | test_cc | test_pr <graph type> ...)\n"
           " [--undirected] [--src=<source_index|randomize|largestdegree>] [--idempotence=<0|1>] [--v]\n"
           " [--iteration-num=<num>] [--traversal-mode=<0|1|2>] [--quick=<0|1>] [--mark-pred]\n"
           " [--queue-sizing=<scale factor>] [--delta-factor=<n>] [--delta=<f>] [--error=<f>] [--max-iter=<n>]\n"
           "Graph types and args:\n"
           "  market <file>                         Matrix-Market coordinate file\n"
           "  rmat --rmat_scale=<s> [--rmat_edgefactor=<f>] [--rmat_seed=<n>]\n");
}

/* reference CompareResults (util/test_utils.cuh:297-340): first mismatch, then CORRECT / INCORRECT */
static int compare_i32(const char *what, const int32_t *got, const int32_t *want, int64_t n)
{
    printf("%s Validity: ", what);
    for (int64_t i = 0; i < n; ++i) {
        if (got[i] != want[i]) {
            printf("\nINCORRECT: [%lu]: %d != %d\n", (unsigned long)i, got[i], want[i]);
            return 1;
        }
    }
    printf("CORRECT\n");
    return 0;
}

int main(int argc, char **argv)
{
    /* invoked under one of the reference's binary names (tests/bfs/test_bfs.cu -> test_bfs, likewise test_dobfs,
     * test_sssp, test_cc, test_pr: `test_bfs market <file> [options]`), the primitive comes from the name */
    char *shifted[256];
    const char *self = strrchr(argv[0], '/');
    self = self ? self + 1 : argv[0];
    if (strncmp(self, "test_", 5) == 0 && argc < 255 &&
        (!strcmp(self + 5, "bfs") || !strcmp(self + 5, "dobfs") || !strcmp(self + 5, "sssp") || !strcmp(self + 5, "cc") ||
         !strcmp(self + 5, "pr"))) {
        shifted[0] = argv[0];
        shifted[1] = (char *)self + 5;
        for (int i = 1; i < argc; ++i) shifted[i + 1] = argv[i];
        ++argc;
        argv = shifted;
    }
    if (argc < 3 || opt_value(argc, argv, "help")) { usage(); return argc < 3; }
    const char *prim = argv[1], *gtype = argv[2];
    const int device = opt_flag(argc, argv, "device", 0);
    const int undirected = opt_flag(argc, argv, "undirected", 0);
    const int quick = opt_flag(argc, argv, "quick", 0);
    const int verbose = opt_flag(argc, argv, "v", 0);
    const int iterations = opt_flag(argc, argv, "iteration-num", 1);
    const int is_dobfs = strcmp(prim, "dobfs") == 0;

    /* ---- graph -------------------------------------------------------------------------- */
    int64_t n = 0, m = 0;
    eid_t *ro = NULL; vid_t *co = NULL; uint32_t *w = NULL;
    const int want_weights = strcmp(prim, "sssp") == 0;
    if (strcmp(gtype, "market") == 0) {
        if (argc < 4) { usage(); return 1; }
        const int sym_read = undirected || is_dobfs;
        /* the reference's cache of the built CSR (market.cuh:254-262: read it if it is there, else build and write it) */
        char cache[1024];
        cache[0] = 0;
        const char *cdir = opt_value(argc, argv, "csr-cache-dir");
        if (cdir && cdir[0]) {
            const char *base = strrchr(argv[3], '/');
            base = base ? base + 1 : argv[3];
            snprintf(cache, sizeof(cache), "%s/.%s_%s_csr", cdir, base, sym_read ? "undirected" : "nonreversed");
        }
        int64_t *cv = NULL;
        if (cache[0] && orc_csr_cache_read(cache, want_weights, &n, &m, &ro, &co, &cv) == 0) {
            printf("  Reading directly from previously stored CSR arrays ...\n");
            if (want_weights) {
                w = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(m > 0 ? m : 1));
                for (int64_t e = 0; e < m; ++e) w[e] = cv ? (uint32_t)cv[e] : 1u;
            }
            orc_free(cv);
        } else {
            int64_t ncoo = 0, *vals = NULL;
            vid_t *rows = NULL, *cols = NULL;
            if (orc_market_read(argv[3], sym_read, 0, want_weights, &n, &ncoo, &rows, &cols, &vals)) {
                fprintf(stderr, "cannot read %s\n", argv[3]);
                return 1;
            }
            ro = (eid_t *)malloc(sizeof(eid_t) * (size_t)(n + 1));
            co = (vid_t *)malloc(sizeof(vid_t) * (size_t)(ncoo > 0 ? ncoo : 1));
            int64_t *vo = want_weights ? (int64_t *)malloc(sizeof(int64_t) * (size_t)(ncoo > 0 ? ncoo : 1)) : NULL;
            m = orc_coo_to_csr(n, ncoo, rows, cols, vals, ro, co, vo);
            if (want_weights) {
                w = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(m > 0 ? m : 1));
                for (int64_t e = 0; e < m; ++e) w[e] = vals ? (uint32_t)vo[e] : 1u;
            }
            if (cache[0] && orc_csr_cache_write(cache, n, m, ro, co, want_weights ? vo : NULL) == 0)
                printf("  Stored the CSR arrays in %s\n", cache);
            orc_free(rows); orc_free(cols); orc_free(vals); free(vo);
        }
    } else if (strcmp(gtype, "rmat") == 0) {
        const int scale = opt_flag(argc, argv, "rmat_scale", 10);
        const int ef = opt_flag(argc, argv, "rmat_edgefactor", 16);
        const uint64_t seed = (uint64_t)opt_flag(argc, argv, "rmat_seed", 42 + scale);
        n = (int64_t)1 << scale;
        const int64_t gen = (int64_t)ef << scale;
        ro = (eid_t *)malloc(sizeof(eid_t) * (size_t)(n + 1));
        co = (vid_t *)malloc(sizeof(vid_t) * (size_t)(2 * gen + 1));
        m = orc_rmat_csr(scale, gen, seed, 1, 0.55, 0.20, 0.20, 0.05, ro, co);    /* rmat.cuh:32-35 defaults, undirected */
        if (want_weights) {
            w = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(m > 0 ? m : 1));
            orc_edge_weights(n, ro, co, seed, w);
        }
    } else { usage(); return 1; }
    const int symmetric = undirected || is_dobfs || strcmp(gtype, "rmat") == 0;
    printf("%s graph: %lld nodes, %lld directed edges%s\n", gtype, (long long)n, (long long)m,
           symmetric ? " (undirected)" : "");
    if (m >= 0x7fffffffLL) { fprintf(stderr, "graph too large for this rig\n"); return 1; }

    /* ---- source (reference: --src=<id> | randomize | largestdegree, test_bfs.cu:530-560) ---- */
    vid_t src = 0;
    const char *sv = opt_value(argc, argv, "src");
    if (sv && strcmp(sv, "randomize") == 0) src = (vid_t)(n ? rand() % n : 0);
    else if (sv && strcmp(sv, "largestdegree") == 0) {
        eid_t best = -1;
        for (int64_t v = 0; v < n; ++v) if (ro[v + 1] - ro[v] > best) { best = ro[v + 1] - ro[v]; src = (vid_t)v; }
        printf("Using highest degree (%lld) vertex: %d\n", (long long)best, src);
    } else if (sv && sv[0]) src = (vid_t)atoi(sv);

    grb_graph_t g = NULL;
    if (grb_graph_from_host(device, n, m, ro, 8, co, w, symmetric, 0, &g)) { fprintf(stderr, "graph upload failed\n"); return 2; }
    int errors = 0;
    grb_stats st;
    memset(&st, 0, sizeof(st));

    if (strcmp(prim, "bfs") == 0 || is_dobfs) {
        grb_bfs_config cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.mark_pred = opt_flag(argc, argv, "mark-pred", 0);
        cfg.idempotence = is_dobfs ? 0 : opt_flag(argc, argv, "idempotence", 0);
        cfg.traversal_mode = opt_flag(argc, argv, "traversal-mode", 0);
        cfg.direction_optimized = is_dobfs;
        const char *qs = opt_value(argc, argv, "queue-sizing");
        cfg.queue_sizing = qs && qs[0] ? atof(qs) : 1.0;
        cfg.record_levels = verbose;
        grb_bfs_t ctx = NULL;
        if (grb_bfs_create(g, &cfg, &ctx)) return 2;
        int32_t *labels = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));
        int32_t *preds = cfg.mark_pred ? (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1)) : NULL;
        for (int it = 0; it < iterations; ++it)
            if (grb_bfs_run(ctx, src, labels, preds, &st)) return 2;
        if (!quick) {
            printf("Computing reference value ...\n");
            int32_t *ref = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));
            double t0 = orc_now_seconds();
            int depth = orc_bfs(n, ro, co, src, ref, NULL);
            printf("CPU BFS finished in %lf msec. cpu_search_depth: %d\n", (orc_now_seconds() - t0) * 1e3, depth);
            /* with idempotence AND predecessors the labels hold predecessors: the reference skips the check
             * (tests/bfs/test_bfs.cu:472-482) */
            if (!(cfg.idempotence && cfg.mark_pred)) errors += compare_i32("Label", labels, ref, n);
            if (preds) {    /* any valid parent: tree edge exists one level up */
                int bad = 0;
                for (int64_t v = 0; v < n && !bad; ++v) {
                    if (ref[v] <= 0) continue;
                    const int32_t p = preds[v];
                    int ok = p >= 0 && p < n && ref[p] + 1 == ref[v];
                    if (ok) { ok = 0; for (eid_t e = ro[p]; e < ro[p + 1]; ++e) if (co[e] == v) { ok = 1; break; } }
                    bad = !ok;
                }
                printf("Predecessor Validity: %s\n", bad ? "INCORRECT" : "CORRECT");
                errors += bad;
            }
            free(ref);
        }
        if (verbose) {
            grb_level_record rec[64];
            int32_t cnt = 0;
            grb_bfs_levels(ctx, rec, 64, &cnt);
            for (int i = 0; i < cnt && i < 64; ++i)
                printf("  level %d kind %d frontier %llu edges %llu discovered %llu  %.3f ms\n", rec[i].level, rec[i].kind,
                       (unsigned long long)rec[i].frontier, (unsigned long long)rec[i].edges,
                       (unsigned long long)rec[i].discovered, rec[i].ms);
        }
        const double rate = st.elapsed_ms > 0 ? (double)st.edges_visited / (st.elapsed_ms * 1000.0) : 0.0;
        printf("[%s] finished. \n elapsed: %.4f ms, rate: %.4f MiEdges/s, search_depth: %lld\n"
               " src: %lld, nodes_visited: %lld, edges_visited: %lld, total queued: %lld\n",
               is_dobfs ? "DOBFS" : "BFS", st.elapsed_ms, rate, (long long)st.iterations, (long long)src,
               (long long)st.nodes_visited, (long long)st.edges_visited, (long long)st.total_queued);
        free(labels); free(preds);
        grb_bfs_destroy(ctx);
    } else if (strcmp(prim, "sssp") == 0) {
        uint32_t *dist = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n + 1));
        int32_t *preds = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));
        const int df = opt_flag(argc, argv, "delta-factor", 1);
        for (int it = 0; it < iterations; ++it)
            if (grb_sssp(g, src, df, 1, dist, preds, &st)) return 2;
        if (!quick) {
            printf("Computing reference value ...\n");
            uint32_t *ref = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n + 1));
            int32_t *rp = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));
            orc_sssp(n, ro, co, w, src, ref, rp);
            errors += compare_i32("Label", (const int32_t *)dist, (const int32_t *)ref, n);
            int bad = 0;
            for (int64_t v = 0; v < n && !bad; ++v) {
                if (ref[v] == 0xffffffffu || v == src) { bad = preds[v] != v; continue; }
                const int32_t p = preds[v];
                int ok = 0;
                if (p >= 0 && p < n) for (eid_t e = ro[p]; e < ro[p + 1]; ++e) if (co[e] == v && ref[p] + w[e] == ref[v]) { ok = 1; break; }
                bad = !ok;
            }
            printf("Predecessor Validity: %s\n", bad ? "INCORRECT" : "CORRECT");
            errors += bad;
            free(ref); free(rp);
        }
        printf("[SSSP] finished. \n elapsed: %.4f ms, iterations: %lld\n src: %lld, nodes_visited: %lld\n",
               st.elapsed_ms, (long long)st.iterations, (long long)src, (long long)st.nodes_visited);
        free(dist); free(preds);
    } else if (strcmp(prim, "cc") == 0) {
        int32_t *labels = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));
        int64_t ncomp = 0;
        for (int it = 0; it < iterations; ++it)
            if (grb_cc(g, labels, &ncomp, &st)) return 2;
        if (!quick) {
            int32_t *ref = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));
            const int64_t want = orc_cc(n, ro, co, ref);
            /* the reference compares the component COUNT only (tests/cc/test_cc.cu:315-320); labels are
             * canonical (min vertex of the component) on both sides here, so they are compared as well */
            if (want == ncomp) printf("CORRECT.\n");
            else { printf("INCORRECT. Ref Component Count: %lld, GPU Computed Component Count: %lld\n", (long long)want, (long long)ncomp); ++errors; }
            errors += compare_i32("Label", labels, ref, n);
            free(ref);
        }
        printf("[CC] finished. \n elapsed: %.4f ms\n num_components: %lld\n", st.elapsed_ms, (long long)ncomp);
        free(labels);
    } else if (strcmp(prim, "pr") == 0) {
        const char *dv = opt_value(argc, argv, "delta"), *ev = opt_value(argc, argv, "error");
        const double delta = dv && dv[0] ? atof(dv) : 0.85, thr = ev && ev[0] ? atof(ev) : 0.01;
        const int max_iter = opt_flag(argc, argv, "max-iter", 20);
        const vid_t psrc = sv ? src : -1;                       /* no --src: global PageRank (pr_problem.cuh:407-413) */
        double *rank = (double *)malloc(sizeof(double) * (size_t)(n + 1));
        for (int it = 0; it < iterations; ++it)
            if (grb_pagerank(g, psrc, delta, thr, max_iter, rank, NULL, NULL, &st)) return 2;
        if (!quick) {
            double *ref = (double *)malloc(sizeof(double) * (size_t)(n + 1));
            orc_pagerank(n, ro, co, delta, thr, (int)st.iterations, psrc, ref);      /* same number of sweeps */
            double worst = 0;
            for (int64_t v = 0; v < n; ++v) { const double d = fabs(rank[v] - ref[v]); if (d > worst) worst = d; }
            printf("Rank Validity: %s (max |diff| %.3e vs fp64 restatement, %lld sweeps)\n",
                   worst <= 1e-6 ? "CORRECT" : "INCORRECT", worst, (long long)st.iterations);
            errors += worst > 1e-6;
            free(ref);
        }
        printf("[PR] finished. \n elapsed: %.4f ms, iterations: %lld\n", st.elapsed_ms, (long long)st.iterations);
        free(rank);
    } else { usage(); return 1; }

    grb_graph_destroy(g);
    free(ro); free(co); free(w);
    printf(errors ? "TEST FAILED\n" : "TEST PASSED\n");
    return errors ? 3 : 0;
}
