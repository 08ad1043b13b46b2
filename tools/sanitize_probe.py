"""Small runs of every primitive for compute-sanitizer (memcheck / racecheck / synccheck), each checked against the CPU oracle:
persistent and host-loop DOBFS, top-down BFS through the streaming advance (TMA ring, mbarriers, harvest), idempotent BFS
(advance + filter), predecessors, the degree-ordered twin, SSSP (near/far split), CC (hooks), PageRank, BC, and the partitioned
BFS / SSSP / CC / PageRank enactors at world 1 (the peer-memory code path with every segment local; both level loops).
usage: compute-sanitizer --tool <tool> python tools/sanitize_probe.py [scale]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gunrockinst_b200 as gr
import oracle
from gunrockinst_b200 import partitioned as pt

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 12
oracle.build()
ro, co = oracle.rmat_csr(scale, 16)
w = oracle.edge_weights(ro, co, 7)
n = ro.size - 1
ok = True


def check(name, cond):
    global ok
    print("%-44s %s" % (name, "ok" if cond else "MISMATCH"), flush=True)
    ok &= bool(cond)


g = gr.Graph.from_csr(ro, co, weights=w, symmetric=True)
want = oracle.bfs(ro, co, 0)
for name, mode in [("bfs persistent dobfs", dict(direction_optimized=True)),
                   ("bfs host loop dobfs", dict(direction_optimized=True, host_loop=True)),
                   ("bfs top-down streaming advance", dict(traversal_mode=1, host_loop=True, stream_min_edges=2)),
                   ("bfs top-down streaming advance, twin", dict(traversal_mode=1, host_loop=True, stream_min_edges=2, degree_order=True)),
                   ("bfs lb advance", dict(traversal_mode=1, host_loop=True, stream_min_edges=1)),
                   ("bfs twc advance", dict(traversal_mode=2)),
                   ("bfs idempotent (advance + filter)", dict(idempotence=True)),
                   ("bfs predecessors", dict(mark_pred=True, direction_optimized=True))]:
    b = gr.Bfs(g, **mode)
    labels, _, _ = b.run(0)
    check(name, np.array_equal(labels, want))
    b.close()
dist, _, _ = gr.sssp(g, 0)
check("sssp near/far", np.array_equal(dist, oracle.sssp(ro, co, w, 0)[0]))
dist, _, _ = gr.sssp(g, 0, delta_factor=16)
check("sssp delta 16", np.array_equal(dist, oracle.sssp(ro, co, w, 0)[0]))
lab, k, _ = gr.cc(g)
wl, wk = oracle.cc(ro, co)
check("cc", np.array_equal(lab, wl) and k == wk)
r, _, st = gr.pagerank(g, max_iter=20)
wr, it = oracle.pagerank(ro, co, 0.85, 1e-6, 20, -1)
check("pagerank (gather)", np.all(np.abs(r - wr) <= 1e-6 * np.maximum(1, np.abs(wr))))
gd = gr.Graph.from_csr(ro, co, symmetric=False)
r, _, st = gr.pagerank(gd, max_iter=20)
check("pagerank (push)", np.all(np.abs(r - wr) <= 1e-6 * np.maximum(1, np.abs(wr))))
b = gr.Bfs(gd, direction_optimized=True)            # transpose built on demand
labels, _, _ = b.run(0)
check("dobfs over a built transpose", np.array_equal(labels, want))
b.close()
gd.close()
bc, _, _ = gr.bc(g, src=0)
check("bc", np.allclose(bc, oracle.bc(ro, co, 0)[0], rtol=1e-4, atol=1e-4))
for persistent in (False, True):
    if persistent:
        os.environ["GRB_MG_PERSISTENT"] = "1"
    eng = pt.DeviceEngine(g)
    os.environ.pop("GRB_MG_PERSISTENT", None)
    for exch in ("peer", "nccl") if not persistent else ("peer",):
        pb = pt.PartitionedBfs(eng, direction_optimized=True, exchange=exch)
        pb.run(0)
        check("partitioned bfs world 1 (%s%s)" % (exch, ", one kernel" if persistent else ""), np.array_equal(pt.gather_labels(eng), want))
    eng.close()
sp = pt.PartitionedSssp(g)
sp.run(0)
check("partitioned sssp world 1", np.array_equal(sp.gather()[0], oracle.sssp(ro, co, w, 0)[0]))
sp.close()
pc = pt.PartitionedCc(g)
pc.run()
check("partitioned cc world 1", np.array_equal(pc.gather()[0], wl))
pc.close()
pp = pt.PartitionedPagerank(g)
pp.run(max_iter=20)
check("partitioned pagerank world 1", np.all(np.abs(pp.gather() - wr) <= 1e-6 * np.maximum(1, np.abs(wr))))
pp.close()
g.close()
print("ALL OK" if ok else "FAILURES", flush=True)
sys.exit(0 if ok else 1)
