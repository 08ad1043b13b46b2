#!/bin/bash
# multi-device check: partitioned parity tests + partitioned bench lines.  usage: bash tools/gpu_multi.sh <ngpus> <tag> [scales...]
n=${1:-2}; tag=${2:-m}; shift; shift
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q > $out/pytest_multi_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_multi_$tag.log
for s in "$@"; do
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $n --steps 16 --warmup 3 --scale $s > $out/bench_${tag}_n${n}_s$s.json 2> $out/bench_${tag}_n${n}_s$s.err
  echo "bench n=$n s=$s rc=$?"; cut -c1-1500 $out/bench_${tag}_n${n}_s$s.json; tail -3 $out/bench_${tag}_n${n}_s$s.err
done
