"""Partitioned CC (grb_mgcc_*) at world size 1 against the single-device CC on the same R-MAT graph.
usage: python tools/mg_cc_probe.py [scale]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gunrockinst_b200 as gr
from gunrockinst_b200 import partitioned as pt

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
g = gr.Graph.rmat(scale)
cc = pt.PartitionedCc(g)
best = None
for _ in range(3):
    st = cc.run()
    if best is None or st["elapsed_ms"] < best["elapsed_ms"]:
        best = st
labels, roots = cc.gather()
single = None
for _ in range(2):
    l1, n1, s1 = gr.cc(g)
    if single is None or s1["elapsed_ms"] < single["elapsed_ms"]:
        single = s1
print(json.dumps(dict(config="mg_cc", scale=scale, n_gpus=1, directed_edges=g.edges, ms=best["elapsed_ms"],
                      iterations=best["iterations"], offers=best["offers"], queued=best["queued"],
                      gpu_launches=best["gpu_launches"], components=roots,
                      single_device=dict(ms=single["elapsed_ms"], iterations=single["iterations"], components=n1,
                                         same_labels=bool(np.array_equal(l1, labels))))), flush=True)
