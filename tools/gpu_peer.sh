#!/bin/bash
# peer-memory exchange check on an N-device box: bash tools/gpu_peer.sh <tag> <scale> <n...>
tag=${1:-p}; scale=${2:-27}; shift; shift
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q > $out/pytest_peer_$tag.log 2>&1; echo "pytest rc=$?"; tail -15 $out/pytest_peer_$tag.log
for n in "$@"; do
 for ex in peer nccl; do
  f=$out/scale_${tag}_${ex}_n${n}_s$scale
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $n --steps 16 --warmup 3 --scale $scale --partitioned --exchange $ex > $f.json 2> $f.err
  echo "n=$n $ex rc=$? lines=$(wc -l < $f.json)"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" $f.err | tail -5
  python - <<P
import json
try:
    d = json.loads([l for l in open("$f.json") if l.startswith("{")][-1])
    x = d["exchange"]
    print("  N=%d %s value=%.1f GTEPS ms=%.3f exch=%.3f comp=%.3f launches=%s setup=%s" % (d["n_gpus"], "$ex", d["value"], d["ms_per_step"], x["ms_per_search"], x["compute_ms_per_search"], d["gpu_launches"], d["setup_s"]))
    for l in x["first_search_levels"]:
        print("    L%d %-4s F=%-9d D=%-9d E=%-10d xB=%-9d exch=%.3f comp=%.3f" % (l["level"], l["kind"], l["frontier"], l["discovered"], l["edges"], l["exchange_bytes"], l["exchange_ms"], l["compute_ms"]))
except Exception as e:
    print("  no line:", e)
P
 done
done
