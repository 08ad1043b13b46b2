#!/bin/bash
# partitioned CC on one box: bash tools/gpu_mgcc.sh <tag> <scale> <n...>   (parity tests first)
tag=${1:-c}; scale=${2:-25}; shift; shift
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q -k "${MGCC_TESTS:-test_cc}" > $out/mgcc_${tag}_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 $out/mgcc_${tag}_pytest.log
for n in "$@"; do
  f=$out/mgcc_${tag}_n${n}_s$scale
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29525 \
    tools/mg_cc_bench.py $scale 3 > $f.json 2> $f.err
  echo "n=$n rc=$?"; cat $f.json | cut -c1-900; grep -v "OMP_NUM_THREADS\|^\*\*\*\*\|NCCL version" $f.err | tail -4
done
