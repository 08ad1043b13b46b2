"""Development probe: per-level trace and timing of BFS variants on one resident graph.
usage: python tools/bfs_probe.py [scale] [runs]   (graph generated on the device when the
library has grb_graph_rmat, else by the oracle on the CPU)"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gunrockinst_b200 as gr

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 20
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 5
t0 = time.time()
gr._capi.lib()
if "grb_graph_rmat" in gr._capi.MISSING or os.environ.get("PROBE_CPU_GRAPH"):
    import oracle
    ro, co = oracle.rmat_csr(scale, 16)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
else:
    g = gr.Graph.rmat(scale)
print("graph: n=%d m=%d built in %.1fs" % (g.nodes, g.edges, time.time() - t0), flush=True)
PEAK = 6539.5
for name, mode in [("warmup", dict(direction_optimized=True)),
                   ("dobfs", dict(direction_optimized=True)), ("dobfs-host", dict(direction_optimized=True, host_loop=1)),
                   ("dobfs-pk", dict(direction_optimized=True, host_loop=2)),
                   ("push-auto", dict()), ("push-host", dict(host_loop=True)), ("push-lb", dict(traversal_mode=1)),
                   ("idempotent", dict(idempotence=True))]:
    if os.environ.get("PROBE_ONLY") and name not in os.environ["PROBE_ONLY"].split(","):
        continue
    bfs = gr.Bfs(g, record_levels=True, **mode)
    best = None
    for r in range(runs):
        _, _, st = bfs.run(0, want_labels=False)
        if best is None or st["elapsed_ms"] < best["elapsed_ms"]:
            best = st
    gteps = best["edges_visited"] / best["elapsed_ms"] / 1e6
    gbs = best["algo_bytes"] / best["elapsed_ms"] / 1e6
    print("%-10s %.3f ms  %.2f GTEPS  algo %.1f GB/s (%.1f%% of %.0f)  launches=%d levels=%d/%d inspected=%d" % (
        name, best["elapsed_ms"], gteps, gbs, 100 * gbs / PEAK, PEAK, best["gpu_launches"],
        best["push_levels"], best["pull_levels"], best["edges_inspected"]), flush=True)
    for lv in bfs.levels() if name != "warmup" else []:
        print("   ", json.dumps(lv))
    bfs.close()
