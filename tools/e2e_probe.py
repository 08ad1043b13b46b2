"""Development probe: (1) wall-clock split of gunrock_bfs_func (GRB_TRACE), pinned vs pageable host CSR,
(2) raw pinned H2D bandwidth of the box, (3) per-level device trace of the persistent DOBFS for a few
bench sources.  usage: python tools/e2e_probe.py [scale]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["GRB_TRACE"] = "1"
import gunrockinst_b200 as gr

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
g = gr.Graph.rmat(scale)
ro, co = g.download()
ro32 = torch.from_numpy(ro.astype(np.int32)).pin_memory()
co32 = torch.from_numpy(co).pin_memory()
dev = torch.empty(co32.numel(), dtype=torch.int32, device="cuda")
for _ in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    dev.copy_(co32, non_blocking=True); torch.cuda.synchronize()
    dt = time.perf_counter() - t0
print("pinned H2D %.1f MB in %.3f ms = %.1f GB/s" % (co32.numel() * 4 / 1e6, dt * 1e3, co32.numel() * 4 / dt / 1e9), flush=True)
del dev
print("-- pinned host CSR", flush=True)
for _ in range(4):
    t0 = time.perf_counter(); gr.gunrock_bfs_func(ro32.numpy(), co32.numpy(), src_node=0)
    print("   call %.3f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
print("-- pageable host CSR", flush=True)
rop, cop = ro.astype(np.int32), co.copy()
for _ in range(3):
    t0 = time.perf_counter(); gr.gunrock_bfs_func(rop, cop, src_node=0)
    print("   call %.3f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)

deg = np.diff(ro)
rng = np.random.RandomState(1234 + scale)
srcs = [0]
while len(srcs) < 6:
    v = int(rng.randint(0, deg.size))
    if deg[v] > 0:
        srcs.append(v)
bfs = gr.Bfs(g, direction_optimized=True, record_levels=True)
bfs.run_many(srcs[:2])
for s in srcs:
    _, _, st = bfs.run(s, want_labels=False)
    print("src %d deg %d: %.3f ms (+reset %.3f)  levels %d  kernel span %d us" % (
        s, deg[s], st["elapsed_ms"], st["reset_ms"], st["iterations"], st["reserved"]))
    for lv in bfs.levels():
        print("    L%d %-8s F=%-9d E=%-10d D=%-9d cand=%-9d %.1f us (switch %.1f)" % (
            lv["level"], lv["kind"], lv["frontier"], lv["edges"], lv["discovered"], lv["candidates"],
            lv["ms"] * 1e3, lv["switch_ms"] * 1e3))
