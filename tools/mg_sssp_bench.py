"""Partitioned SSSP (grb_mgsssp_*) on R-MAT with symmetric weights 1..64 under torchrun:
   python -m torch.distributed.run --nproc-per-node N tools/mg_sssp_bench.py [scale] [runs]
One JSON line on rank 0: time per search (CUDA events inside the library, max over ranks), iterations,
relaxations, and a whole-result check against the single-device near/far SSSP (grb_sssp, itself
oracle-checked by tools/run_configs.py c3) through two size-independent sums."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gunrockinst_b200 as gr
from gunrockinst_b200 import partitioned as pt

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 24
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 5
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", rank))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29541")
sys.stdout.flush()
saved = os.dup(1)
os.dup2(2, 1)
torch.cuda.set_device(local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
t0 = time.time()
lp = world.bit_length() - 1
g = gr.Graph.rmat(scale, 16, device=local, log_parts=lp, rank=rank, weight_seed=42 + scale)
sp = pt.PartitionedSssp(g)
cnt = torch.tensor([g.edges], dtype=torch.int64, device=dev)
dist.all_reduce(cnt)
best = None
for r in range(runs + 1):
    dist.barrier()
    st = sp.run(0)
    t = torch.tensor([st["elapsed_ms"]], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    st["elapsed_ms"] = float(t.item())
    if r > 0 and (best is None or st["elapsed_ms"] < best["elapsed_ms"]):
        best = st
d, p = sp.result()
reached = d != 0xFFFFFFFF
sums = torch.tensor([int(d[reached].astype(np.int64).sum()), int(reached.sum())], dtype=torch.int64, device=dev)
dist.all_reduce(sums)
line = None
if rank == 0:
    check = None
    try:        # the same graph on one device, near/far SSSP
        g1 = gr.Graph.rmat(scale, 16, device=local, weight_seed=42 + scale)
        d1, _, st1 = gr.sssp(g1, 0)
        r1 = d1 != 0xFFFFFFFF
        check = dict(match=bool(int(d1[r1].astype(np.int64).sum()) == int(sums[0]) and int(r1.sum()) == int(sums[1])),
                     single_device_ms=st1["elapsed_ms"], single_device_iterations=st1["iterations"])
        g1.close()
    except Exception as e:      # not enough memory next to the partitioned slice, ...
        check = dict(match=None, error=str(e)[:200])
    m = int(cnt.item())
    line = dict(config="mg_sssp", scale=scale, n_gpus=world, directed_edges=m, ms=best["elapsed_ms"],
                iterations=best["iterations"], relaxations=best["relaxations"], queued=best["queued"],
                gteps=m / best["elapsed_ms"] / 1e6, grelax_per_s=best["relaxations"] / best["elapsed_ms"] / 1e6,
                reached=int(sums[1]), dist_sum=int(sums[0]), check_vs_single_device=check,
                exchange="peer memory: 64-bit atomicMin over NVLink from the advance kernel",
                setup_s=round(time.time() - t0, 1))
sp.close()
g.close()
dist.destroy_process_group()
sys.stdout.flush()
os.dup2(saved, 1)
if line:
    print(json.dumps(line), flush=True)
