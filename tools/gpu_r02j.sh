#!/bin/bash
# 1-GPU session for the row-binned passes (oprtr/row_bins.cuh): A/B against the LB routes with parity, the whole GPU
# suite on the new defaults, the bench line, full ncu captures of the two new kernels.  Cheapest and most telling first.
tag=${1:-r02j}
out=gpurun_out; mkdir -p $out
timeout 120 python tools/rows_ab.py 16 24 1 > $out/rows_ab_s16_$tag.jsonl 2> $out/rows_ab_s16_$tag.err; echo "ab16 rc=$?"; cat $out/rows_ab_s16_$tag.jsonl; tail -3 $out/rows_ab_s16_$tag.err
timeout 200 python tools/rows_ab.py 25 24 2 > $out/rows_ab_s25_$tag.jsonl 2> $out/rows_ab_s25_$tag.err; echo "ab25 rc=$?"; cat $out/rows_ab_s25_$tag.jsonl; tail -3 $out/rows_ab_s25_$tag.err
timeout 600 python -m pytest tests -m gpu -q --timeout 240 > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -15 $out/pytest_gpu_$tag.log | cut -c1-300
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke_$tag.log 2>&1; echo "smoke rc=$?"; tail -1 $out/smoke_$tag.log
timeout 600 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"; cut -c1-300 $out/bench_$tag.json; tail -3 $out/bench_$tag.err
TAG=$tag python - <<'P'
import json, os
try:
    d = json.load(open("gpurun_out/bench_%s.json" % os.environ["TAG"]))
    print(json.dumps({k: d["configs"].get(k) for k in ("c4_cc_s25", "c4_pr_s25")})[:1500])
except Exception as e:
    print("no configs", e)
P
timeout 240 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'PrRowSweepTicketKernel|CcRowHookKernel|CcFlattenKernel' -c 5 -f -o $out/prof_rows_$tag \
    python tools/rows_ab.py 25 2 1 > $out/ncu_rows_$tag.log 2>&1
echo "rows capture rc=$?"; tail -2 $out/ncu_rows_$tag.log
