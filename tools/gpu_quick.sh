#!/bin/bash
# Quick GPU check: parity tests, per-level probe, short bench.  usage: bash tools/gpu_quick.sh <tag> [scale]
tag=${1:-q}; scale=${2:-22}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest_gpu_$tag.log
timeout 600 python tools/e2e_probe.py $scale > $out/probe_$tag.log 2>&1; echo "probe rc=$?"; grep -v "^\[BFS\]\|^ elapsed\|^ src\|^ nodes_v\|^ edges_v\|^ rate\|^ search_d" $out/probe_$tag.log
timeout 600 python bench.py --no-cpu --e2e-steps 4 > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"; cat $out/bench_$tag.json; tail -3 $out/bench_$tag.err
