#!/bin/bash
# partitioned SSSP table on one box: bash tools/gpu_mgsssp.sh <tag> <scale> <n...>
tag=${1:-s}; scale=${2:-24}; shift; shift
out=gpurun_out; mkdir -p $out
for n in "$@"; do
  f=$out/mgsssp_${tag}_n${n}_s$scale
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29519 \
    tools/mg_sssp_bench.py $scale 5 > $f.json 2> $f.err
  echo "n=$n rc=$?"; cat $f.json | cut -c1-700; grep -v "OMP_NUM_THREADS\|^\*\*\*\*\|NCCL version" $f.err | tail -4
done
