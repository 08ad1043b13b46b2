"""BASELINE.json configs at full size on one B200, each checked against the CPU oracle.

usage: python tools/run_configs.py [c2] [bfs25] [c3] [c4cc] [c4pr] [--no-check] [--out FILE]

  c2     direction-optimising BFS, R-MAT scale 22, source 0 + 16 seeded sources
  bfs25  the same at scale 25 (the north star's single-GPU target)
  c3     SSSP, R-MAT scale 24, symmetric integer weights 1..64, source 0
  c4cc   connected components, R-MAT scale 25
  c4pr   PageRank (delta 0.85, threshold 1e-6), R-MAT scale 25

Every line printed is one JSON object: device time (CUDA events inside the library), throughput,
algorithmic bytes over that time, and `parity` = bit-exact comparison with
the oracle on the identical CSR (PageRank: max |diff| against the fp64 restatement run for the
same number of sweeps).  The oracle is test infrastructure: it is only used here as the checker.
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gunrockinst_b200 as gr

PEAK = 6548.5
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def emit(out, rec):
    line = json.dumps(rec)
    print(line, flush=True)
    if out:
        with open(out, "a") as f:
            f.write(line + "\n")


def best_of(fn, runs):
    best = None
    for _ in range(runs):
        r = fn()
        if best is None or r[-1]["elapsed_ms"] < best[-1]["elapsed_ms"]:
            best = r
    return best


def pick_sources(ro, count, seed):
    deg = np.diff(ro)
    rng = np.random.RandomState(seed)
    cand = np.nonzero(deg > 0)[0]
    return [0] + rng.choice(cand, count, replace=False).tolist()


def run_bfs(scale, check, out, tag):
    import oracle
    t0 = time.time()
    g = gr.Graph.rmat(scale)
    ro, co = g.download()
    sources = pick_sources(ro, 16, 1234 + scale)
    bfs = gr.Bfs(g, direction_optimized=True)
    bfs.run_many(sources[:3])
    stats = bfs.run_many(sources)
    ms = [s["elapsed_ms"] + s["reset_ms"] for s in stats]
    teps = [s["edges_visited"] / (m * 1e-3) / 1e9 for s, m in zip(stats, ms)]
    hmean = len(teps) / sum(1.0 / t for t in teps)
    parity = None
    if check:
        parity = True
        for src in sources[:3]:
            labels, _, _ = bfs.run(src)
            parity = parity and bool(np.array_equal(labels, oracle.bfs(ro, co, src)))
    # top-down only (the reference's bfs_enactor structure) for the push roofline
    td = gr.Bfs(g)
    td.run(0, want_labels=False)
    _, _, st = td.run(0, want_labels=False)
    emit(out, {"config": tag, "scale": scale, "nodes": g.nodes, "directed_edges": g.edges,
               "sources": len(sources), "gteps_harmonic_mean": hmean, "gteps_src0": teps[0],
               "ms_src0": ms[0], "ms_mean": float(np.mean(ms)), "levels_src0": stats[0]["iterations"],
               "algo_gbs_src0": stats[0]["algo_bytes"] / (ms[0] * 1e-3) / 1e9,
               "topdown_ms_src0": st["elapsed_ms"], "topdown_gteps_src0": st["edges_visited"] / st["elapsed_ms"] / 1e6,
               "topdown_algo_gbs": st["algo_bytes"] / st["elapsed_ms"] / 1e6,
               "topdown_frac_of_hbm_peak": st["algo_bytes"] / st["elapsed_ms"] / 1e6 / PEAK,
               "parity": parity, "setup_s": round(time.time() - t0, 1)})
    bfs.close(); td.close(); g.close()


def run_sssp(scale, check, out):
    import oracle
    t0 = time.time()
    g = gr.Graph.rmat(scale, weight_seed=7)
    ro, co, w = g.download(want_weights=True)
    gr.sssp(g, 0)
    dist, preds, st = best_of(lambda: gr.sssp(g, 0), 3)
    parity = None
    if check:
        want, _ = oracle.sssp(ro, co, w, 0)
        parity = bool(np.array_equal(dist, want))
    reached = dist != np.uint32(0xFFFFFFFF)
    emit(out, {"config": "c3_sssp", "scale": scale, "nodes": g.nodes, "directed_edges": g.edges,
               "ms": st["elapsed_ms"], "iterations": st["iterations"], "relaxations": st["edges_inspected"],
               "gteps": int(np.diff(ro)[reached].sum()) / st["elapsed_ms"] / 1e6,
               "algo_gbs": st["algo_bytes"] / st["elapsed_ms"] / 1e6,
               "frac_of_hbm_peak": st["algo_bytes"] / st["elapsed_ms"] / 1e6 / PEAK,
               "parity": parity, "weights_symmetric": True, "setup_s": round(time.time() - t0, 1)})
    g.close()


def run_cc(scale, check, out):
    import oracle
    t0 = time.time()
    g = gr.Graph.rmat(scale)
    gr.cc(g)
    labels, k, st = best_of(lambda: gr.cc(g), 3)
    parity = None
    if check:
        ro, co = g.download()
        want, ncomp = oracle.cc(ro, co)
        parity = bool(np.array_equal(labels, want) and k == ncomp)
    emit(out, {"config": "c4_cc", "scale": scale, "nodes": g.nodes, "directed_edges": g.edges,
               "ms": st["elapsed_ms"], "components": k, "launches": st["gpu_launches"],
               "algo_gbs": st["algo_bytes"] / st["elapsed_ms"] / 1e6,
               "frac_of_hbm_peak": st["algo_bytes"] / st["elapsed_ms"] / 1e6 / PEAK,
               "parity": parity, "setup_s": round(time.time() - t0, 1)})
    g.close()


def run_pr(scale, check, out, check_iters=5):
    import oracle
    t0 = time.time()
    g = gr.Graph.rmat(scale)
    gr.pagerank(g, max_iter=2)
    ranks, _, st = gr.pagerank(g, delta=0.85, threshold=1e-6, max_iter=1000)
    parity = maxdiff = None
    if check:
        ro, co = g.download()
        got, _, _ = gr.pagerank(g, delta=0.85, threshold=1e-6, max_iter=check_iters)
        want = oracle.pagerank(ro, co, delta=0.85, threshold=1e-6, max_iter=check_iters)
        want = want[0] if isinstance(want, tuple) else want
        maxdiff = float(np.max(np.abs(got - want)))
        parity = bool(maxdiff <= 1e-6)
    per_iter = st["elapsed_ms"] / max(st["iterations"], 1)
    emit(out, {"config": "c4_pr", "scale": scale, "nodes": g.nodes, "directed_edges": g.edges,
               "ms": st["elapsed_ms"], "iterations": st["iterations"], "ms_per_iteration": per_iter,
               "algo_gbs": st["algo_bytes"] / st["elapsed_ms"] / 1e6,
               "frac_of_hbm_peak": st["algo_bytes"] / st["elapsed_ms"] / 1e6 / PEAK,
               "parity_after_%d_sweeps" % check_iters: parity, "max_abs_diff": maxdiff,
               "rank_sum": float(ranks.sum()), "setup_s": round(time.time() - t0, 1)})
    g.close()


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    check = "--no-check" not in sys.argv
    out = None
    if "--out" in sys.argv:
        out = sys.argv[sys.argv.index("--out") + 1]
        args = [a for a in args if a != out]
    if not args:
        args = ["c2", "bfs25", "c3", "c4cc", "c4pr"]
    gr._capi.lib()
    for a in args:
        if a == "c2":
            run_bfs(22, check, out, "c2_dobfs")
        elif a == "bfs25":
            run_bfs(25, check, out, "dobfs_s25")
        elif a == "c3":
            run_sssp(24, check, out)
        elif a == "c4cc":
            run_cc(25, check, out)
        elif a == "c4pr":
            run_pr(25, check, out)
        elif a.startswith("bfs"):
            run_bfs(int(a[3:]), check, out, "dobfs_s" + a[3:])
        else:
            raise SystemExit("unknown config " + a)


if __name__ == "__main__":
    main()
