#!/bin/bash
# short closing check of the CTA-aggregated row binning: its contract test, the CC and PageRank suites, A/B at scale 25
tag=${1:-r02n}
out=gpurun_out; mkdir -p $out
timeout 100 python -m pytest tests/test_operators_gpu.py tests/test_cc_gpu.py tests/test_pr_gpu.py -m gpu -q -k "not scale22" --timeout 60 > $out/pytest_rows_$tag.log 2>&1; echo "pytest rc=$?"; tail -4 $out/pytest_rows_$tag.log | cut -c1-300
timeout 60 python tools/rows_ab.py 25 8 1 > $out/rows_ab_s25_$tag.jsonl 2> $out/rows_ab_s25_$tag.err; echo "ab25 rc=$?"; cat $out/rows_ab_s25_$tag.jsonl; tail -3 $out/rows_ab_s25_$tag.err
