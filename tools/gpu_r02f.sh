#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 300 python -m pytest tests/test_graph_gpu.py tests/test_operators_gpu.py -m gpu -x -q > $out/pytest_f.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_f.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 1 --steps 16 --warmup 3 --scale 24 --no-cpu > $out/bench_f.json 2> $out/bench_f.err; echo "rc=$?"; tail -2 $out/bench_f.err
python - <<PY
import json
r=json.loads(open("$out/bench_f.json").read().strip().splitlines()[-1])
print(r["value"], r["e2e"], r["e2e_from_host"])
PY
