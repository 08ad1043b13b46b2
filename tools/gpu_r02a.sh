#!/bin/bash
# round-2 first box session: parity tests, streaming-kernel probe at scale 25 and 22, bench without the full-size configs
tag=${1:-r02a}
out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $out/smi_$tag.log
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest_gpu_$tag.log
timeout 600 python tools/stream_eval.py 25 3 > $out/stream_eval_s25_$tag.jsonl 2> $out/stream_eval_s25_$tag.err; echo "eval25 rc=$?"
timeout 300 python tools/stream_eval.py 22 3 > $out/stream_eval_s22_$tag.jsonl 2> $out/stream_eval_s22_$tag.err; echo "eval22 rc=$?"
timeout 900 python bench.py --no-configs > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"; tail -3 $out/bench_$tag.err
