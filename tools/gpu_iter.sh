#!/bin/bash
# quick iteration on the streaming advance: its parity tests, then the probe at the given scales
# usage: bash tools/gpu_iter.sh <tag> "<variants>" <scale> [<scale> ...]
tag=$1; variants=$2; shift 2
out=gpurun_out; mkdir -p $out
timeout 150 python -m pytest tests/test_bfs_gpu.py -m gpu -x -q -k "stream or degree_order or idem or all_modes or fixtures" > $out/pytest_iter_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_iter_$tag.log
for s in "$@"; do
  timeout 120 python tools/stream_eval.py $s 3 $variants > $out/stream_eval_s${s}_$tag.jsonl 2> $out/stream_eval_s${s}_$tag.err; echo "eval$s rc=$?"
  python - <<PY
import json
for l in open("$out/stream_eval_s${s}_$tag.jsonl"):
    r=json.loads(l)
    if "variant" not in r: continue
    print("%-18s ms=%.3f gteps=%.0f frac=%.3f same=%s"%(r["variant"],r["ms"],r["gteps"],r["frac"],r["labels_equal_first_variant"]), [ (lv["kind"],lv["edges"],round(lv["ms"],3)) for lv in r["level_records"] if lv["ms"]>0.05])
PY
done
