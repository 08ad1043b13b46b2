#!/bin/bash
# final multi-device session on an 8-GPU box: partitioned parity tests, scale-27 BFS table (peer exchange),
# scale-24 SSSP table, reference arm under torchrun
tag=${1:-fin}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q > $out/pytest_multi_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_multi_$tag.log
for n in 8 4 2 1; do
  f=$out/scale_${tag}_peer_n${n}_s27
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $n --steps 32 --warmup 3 --scale 27 --partitioned > $f.json 2> $f.err
  echo "bfs n=$n rc=$? lines=$(wc -l < $f.json)"; grep -v "OMP_NUM_THREADS\|^\*\*\*\*\|NCCL version" $f.err | tail -3
  python - <<P
import json
try:
    d = json.loads(open("$f.json").read()); x = d["exchange"]
    print("  N=%d value=%.1f GTEPS ms=%.3f exch=%.3f comp=%.3f launches=%s roofline=%s" % (d["n_gpus"], d["value"], d["ms_per_step"], x["ms_per_search"], x["compute_ms_per_search"], d["gpu_launches"], d["roofline"]["frac"]))
except Exception as e:
    print("  no line:", e)
P
done
bash tools/gpu_mgsssp.sh $tag 24 8 4 2 1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 \
  bench.py --impl reference --gpus 2 --steps 4 --warmup 1 > $out/ref_n2_$tag.json 2> $out/ref_n2_$tag.err
echo "ref n=2 rc=$? lines=$(wc -l < $out/ref_n2_$tag.json)"; cut -c1-300 $out/ref_n2_$tag.json; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" $out/ref_n2_$tag.err | tail -3
