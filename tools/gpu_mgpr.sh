#!/bin/bash
# partitioned PageRank on one box: bash tools/gpu_mgpr.sh <tag> <scale> <n...>   (parity tests first)
tag=${1:-p}; scale=${2:-24}; shift; shift
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q -k "${MGPR_TESTS:-pagerank}" > $out/mgpr_${tag}_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 $out/mgpr_${tag}_pytest.log
for n in "$@"; do
  f=$out/mgpr_${tag}_n${n}_s$scale
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29523 \
    tools/mg_pr_bench.py $scale 2 > $f.json 2> $f.err
  echo "n=$n rc=$?"; cat $f.json | cut -c1-900; grep -v "OMP_NUM_THREADS\|^\*\*\*\*\|NCCL version" $f.err | tail -4
done
