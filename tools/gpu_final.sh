#!/bin/bash
# last 1-GPU session of a round: whole GPU suite (no -x: every failure is listed), the bench line, the
# single-device PageRank timing at scale 25 through the partitioned bench's cross-check
tag=${1:-final}
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests -m gpu -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -4 $out/pytest_gpu_$tag.log
timeout 600 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"; cut -c1-1500 $out/bench_$tag.json; tail -3 $out/bench_$tag.err
timeout 300 python tools/mg_pr_bench.py 25 1 > $out/mgpr_${tag}_n1_s25.json 2> $out/mgpr_${tag}_n1_s25.err; echo "pr rc=$?"; cut -c1-900 $out/mgpr_${tag}_n1_s25.json
