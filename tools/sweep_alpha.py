"""Development probe: direction-switch parameters (Beamer's alpha / beta) against whole-search time
over the bench's 64 sources.  usage: python tools/sweep_alpha.py [scale]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gunrockinst_b200 as gr

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
g = gr.Graph.rmat(scale)
ro, _ = g.download()
deg = np.diff(ro)
rng = np.random.RandomState(1234 + scale)
srcs = [0]
while len(srcs) < 64:
    v = int(rng.randint(0, deg.size))
    if deg[v] > 0:
        srcs.append(v)
for alpha in (3, 5, 8, 11, 14, 20, 40):
    for beta in (8, 24, 72):
        bfs = gr.Bfs(g, direction_optimized=True, alpha=alpha, beta=beta)
        bfs.run_many(srcs[:4])
        st = bfs.run_many(srcs)
        ms = sum(s["elapsed_ms"] + s["reset_ms"] for s in st)
        ev = sum(s["edges_visited"] for s in st)
        print("alpha %5.1f beta %5.1f : %.4f ms/search  %.1f GTEPS  inspected/search %.2fM  levels %.2f" % (
            alpha, beta, ms / len(st), ev / ms / 1e6, sum(s["edges_inspected"] for s in st) / len(st) / 1e6,
            sum(s["iterations"] for s in st) / len(st)), flush=True)
        bfs.close()
