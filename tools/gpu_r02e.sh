#!/bin/bash
# persistent partitioned BFS on ONE device: parity at world 1 (peer exchange path), then the torchrun N=1 line at scale 24,
# with the host loop beside it
out=gpurun_out; mkdir -p $out
export GRB_PEER_TIMEOUT_MS=3000
timeout 200 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q -k "world1" > $out/pytest_mgp1.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_mgp1.log
for mode in host; do
  true
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 1 --steps 32 --warmup 3 --scale 24 --no-cpu > $out/bench_mgp1_$mode.json 2> $out/bench_mgp1_$mode.err; echo "$mode rc=$?"; tail -2 $out/bench_mgp1_$mode.err
  python - <<PY
import json
r=json.loads(open("$out/bench_mgp1_$mode.json").read().strip().splitlines()[-1])
print("$mode", {k:r[k] for k in ("value","ms_per_step","gpu_launches")}, r["parity"]["ok"], r["single_device_same_graph"]["gteps"], r["exchange"]["ms_per_search"], r["exchange"]["compute_ms_per_search"])
for l in r["exchange"]["first_search_levels"]: print("   ", l["level"], l["kind"], l["frontier"], l["edges"], round(l["compute_ms"],4), round(l["exchange_ms"],4))
PY
done
