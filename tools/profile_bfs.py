"""One DOBFS run and one push-only (LB) run on a device-built R-MAT graph: the command
that is put under ncu.  usage: python tools/profile_bfs.py [scale]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gunrockinst_b200 as gr

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
g = gr.Graph.rmat(scale)
for mode in (dict(direction_optimized=True), dict(traversal_mode=1)):
    bfs = gr.Bfs(g, **mode)
    for _ in range(2):
        _, _, st = bfs.run(0, want_labels=False)
    print(mode, "%.3f ms" % st["elapsed_ms"], st["edges_visited"] / st["elapsed_ms"] / 1e6, "GTEPS")
    bfs.close()
