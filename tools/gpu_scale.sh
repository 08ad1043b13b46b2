#!/bin/bash
# Scaling table of the partitioned BFS on ONE box: bash tools/gpu_scale.sh <tag> <scale> <n...>
tag=${1:-sc}; scale=${2:-27}; shift; shift
out=gpurun_out; mkdir -p $out
for n in "$@"; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $n --steps 16 --warmup 3 --scale $scale --partitioned > $out/scale_${tag}_n${n}_s$scale.json 2> $out/scale_${tag}_n${n}_s$scale.err
  echo "n=$n rc=$?"; cut -c1-400 $out/scale_${tag}_n${n}_s$scale.json; tail -3 $out/scale_${tag}_n${n}_s$scale.err
  python - <<P
import json
try:
    d = json.load(open("$out/scale_${tag}_n${n}_s$scale.json"))
    print("  N=%d value=%.1f GTEPS ms=%.3f exch=%.3f comp=%.3f launches=%s setup=%s" % (d["n_gpus"], d["value"], d["ms_per_step"], d["exchange"]["ms_per_search"], d["exchange"]["compute_ms_per_search"], d["gpu_launches"], d["setup_s"]))
except Exception as e:
    print("  no line:", e)
P
done
