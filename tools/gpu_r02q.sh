#!/bin/bash
# the CC hook pass with tickets: CC suite, A/B at scale 16 (oracle) and 25
out=gpurun_out; mkdir -p $out
timeout 40 python -m pytest tests/test_cc_gpu.py -m gpu -q --timeout 30 > $out/pytest_cc_r02q.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_cc_r02q.log | cut -c1-300
timeout 20 python tools/rows_ab.py 16 2 2 cc > $out/rows_ab_cc_s16_r02q.jsonl 2>&1; cat $out/rows_ab_cc_s16_r02q.jsonl | cut -c1-700
timeout 30 python tools/rows_ab.py 25 2 3 cc > $out/rows_ab_cc_s25_r02q.jsonl 2>&1; cat $out/rows_ab_cc_s25_r02q.jsonl | cut -c1-700
