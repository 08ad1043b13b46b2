#!/bin/bash
# the scaling-run entry on one device: torchrun N=1 must run the partitioned workload with its reference point and parity
out=gpurun_out; mkdir -p $out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 1 --steps 16 --warmup 3 --scale 24 > $out/bench_torchrun1_s24.json 2> $out/bench_torchrun1_s24.err; echo "rc=$?"; tail -3 $out/bench_torchrun1_s24.err
python - <<PY
import json
r=json.loads(open("$out/bench_torchrun1_s24.json").read().strip().splitlines()[-1])
print({k:r[k] for k in ("value","ms_per_step","n_gpus","single_device_same_graph","speedup_vs_single_device","e2e","gpu_launches")})
print(r["parity"]); print(r["config"]["workload"]); print(r["exchange"]["ms_per_search"], r["exchange"]["compute_ms_per_search"], r["exchange"]["levels_per_search"])
PY
