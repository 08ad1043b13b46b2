"""Wall-clock split of gunrock_bfs_func calls (GRB_TRACE=1) on a scale-S R-MAT graph in pinned host memory.
usage: python tools/e2e_trace.py [scale] [calls]"""
import os
import sys
import time

os.environ["GRB_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gunrockinst_b200 as gr

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 22
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 6
g = gr.Graph.rmat(scale)
ro, co = g.download()
g.close()
ro32 = torch.from_numpy(ro.astype(np.int32)).pin_memory()
co32 = torch.from_numpy(co).pin_memory()
for i in range(calls):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    gr.gunrock_bfs_func(ro32.numpy(), co32.numpy(), src_node=0)
    print("call %d: %.3f ms" % (i, 1e3 * (time.perf_counter() - t0)), file=sys.stderr, flush=True)
