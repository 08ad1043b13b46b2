"""Development probe (round 2): BFS variants on one resident R-MAT graph -- the round-1 kernels, the
streaming advance (lb_stream.cuh), the degree-ordered twin -- with their per-level records and a
labels-equal check between all variants.
usage: python tools/stream_eval.py <scale> [runs] [variant,variant,...]   -> JSON lines on stdout"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gunrockinst_b200 as gr
if os.environ.get("GRB_EVAL_LIB"):          # A/B runs: another build of the library (before the first call loads it)
    gr._capi.LIB_PATH = os.environ["GRB_EVAL_LIB"]

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 3
only = sys.argv[3].split(",") if len(sys.argv) > 3 else None
PEAK = 6531.3
t0 = time.time()
g = gr.Graph.rmat(scale)
print(json.dumps({"graph": {"scale": scale, "n": g.nodes, "m": g.edges, "build_s": round(time.time() - t0, 1)}}), flush=True)

VARIANTS = [
    ("td_r1", dict(traversal_mode=1, host_loop=1, stream_min_edges=1)),
    ("td_stream", dict(traversal_mode=1, host_loop=1, stream_min_edges=1 << 20)),
    ("td_stream_twin", dict(traversal_mode=1, host_loop=1, degree_order=True)),
    ("td_r1_twin", dict(traversal_mode=1, host_loop=1, stream_min_edges=1, degree_order=True)),
    ("dobfs_r1", dict(direction_optimized=True, stream_min_edges=1)),
    ("dobfs_twin", dict(direction_optimized=True, degree_order=True)),
    ("dobfs_host_r1", dict(direction_optimized=True, host_loop=1, stream_min_edges=1)),
    ("dobfs_host_twin", dict(direction_optimized=True, host_loop=1, degree_order=True)),
    ("idem_r1", dict(idempotence=True, stream_min_edges=1)),
    ("idem_stream", dict(idempotence=True, stream_min_edges=1 << 20)),
    ("idem_stream_twin", dict(idempotence=True, degree_order=True)),
]
ref_labels = None
for name, mode in VARIANTS:
    if only and name not in only:
        continue
    t1 = time.time()
    bfs = gr.Bfs(g, record_levels=True, **mode)
    labels, _, st = bfs.run(0)
    setup_s = time.time() - t1
    if ref_labels is None:
        ref_labels = labels
    same = bool(np.array_equal(labels, ref_labels))
    best, best_lv = None, None
    for r in range(runs):
        _, _, st = bfs.run(0, want_labels=False)
        if best is None or st["elapsed_ms"] < best["elapsed_ms"]:
            best, best_lv = st, bfs.levels()
    ms = best["elapsed_ms"]
    rec = {"variant": name, "mode": mode, "ms": ms, "gteps": best["edges_visited"] / ms / 1e6,
           "algo_gbs": best["algo_bytes"] / ms / 1e6, "frac": best["algo_bytes"] / ms / 1e6 / PEAK,
           "launches": best["gpu_launches"], "levels": [best["push_levels"], best["pull_levels"]],
           "inspected": best["edges_inspected"], "labels_equal_first_variant": same, "first_run_s": round(setup_s, 2),
           "level_records": [{k: (round(v, 4) if isinstance(v, float) else v) for k, v in lv.items()} for lv in best_lv]}
    print(json.dumps(rec), flush=True)
    bfs.close()
g.close()
