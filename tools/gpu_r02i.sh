#!/bin/bash
# tail kernel at N devices: parity (BFS world tests), then the scale-27 line with and without it
n=${1:-2}
out=gpurun_out; mkdir -p $out
export GRB_PEER_TIMEOUT_MS=8000
timeout 600 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q -k "test_multi_device_matches_oracle" > $out/pytest_tail_n$n.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_tail_n$n.log
for tail in 1 0; do
  GRB_MG_TAIL=$tail timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$tail bench.py --gpus $n --steps 64 --warmup 3 --no-single-ref --no-e2e-from-host > $out/tail${tail}_n$n.json 2> $out/tail${tail}_n$n.err; echo "tail=$tail rc=$?"
  python - <<PY
import json
r=json.loads(open("$out/tail${tail}_n$n.json").read().strip().splitlines()[-1])
print("tail=$tail n=$n gteps=%.0f ms=%.3f launches/search/gpu=%.1f exch=%.3f comp=%.3f"%(r["value"], r["ms_per_step"], r["gpu_launches"]/r["steps"]/r["n_gpus"], r["exchange"]["ms_per_search"], r["exchange"]["compute_ms_per_search"]))
for l in r["exchange"]["first_search_levels"]: print("   ", l["level"], l["kind"], l["frontier"], l["edges"], "comp %.4f exch %.4f"%(l["compute_ms"], l["exchange_ms"]))
PY
done
