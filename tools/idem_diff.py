import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np
import gunrockinst_b200 as gr
g = gr.Graph.rmat(int(sys.argv[1]))
a = gr.Bfs(g, idempotence=True, stream_min_edges=1)
la, _, sa = a.run(0)
for twin in (False, True):
    b = gr.Bfs(g, idempotence=True, degree_order=twin, record_levels=True)
    lb, _, sb = b.run(0)
    bad = np.nonzero(la != lb)[0]
    print("twin", twin, "mismatches", bad.size, "edges_visited", sa["edges_visited"], sb["edges_visited"])
    if bad.size:
        print(" first", bad[:10], la[bad[:10]], lb[bad[:10]])
        print(" depth histogram of the r1 labels of mismatching vertices", np.bincount(la[bad] + 1))
        print(" stream labels there", np.unique(lb[bad], return_counts=True))
        print(" id range", bad.min(), bad.max(), "hot(<1.2M)", int((bad < 1200000).sum()))
    print([ (r["kind"], r["frontier"], r["discovered"]) for r in b.levels()])
    b.close()
