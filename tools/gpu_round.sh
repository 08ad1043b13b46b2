#!/bin/bash
# One GPU-box session: parity tests, bench (both arms), launch list, full ncu capture.
# usage (from the repo root, through gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_gpu_$tag.log
timeout 900 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"; cat $out/bench_$tag.json; tail -5 $out/bench_$tag.err
timeout 900 python bench.py --impl reference --steps 32 --warmup 3 > $out/bench_ref_$tag.json 2> $out/bench_ref_$tag.err; echo "ref rc=$?"; cat $out/bench_ref_$tag.json
small="python bench.py --steps 2 --warmup 1 --no-cpu --e2e-steps 1"
$small > $out/plain_small.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/launches_$tag.csv $small > $out/ncu_launches.log 2>&1
echo "launch list rc=$?"
probe="python tools/bfs_probe.py 22 1"
$probe > $out/plain_probe.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'pull|Expand' -c 8 -f -o $out/prof_$tag $probe > $out/ncu_full.log 2>&1
echo "full capture rc=$?"; tail -3 $out/ncu_full.log
