#!/bin/bash
# A/B of pull-sweep builds: the DOBFS variants of tools/stream_eval.py with each library build
# gunrockinst_b200/lib/libgunrock_<name>.so ("main" = the product library)
# usage: bash tools/gpu_pull_ab.sh <tag> "<scales>" <name> [<name> ...]
tag=$1; scales=$2; shift 2
out=gpurun_out; mkdir -p $out
for s in $scales; do
  for rep in 1 2; do
    for name in "$@"; do
      if [ $name = main ]; then unset GRB_EVAL_LIB; else export GRB_EVAL_LIB=$PWD/gunrockinst_b200/lib/libgunrock_$name.so; fi
      timeout 200 python tools/stream_eval.py $s 5 dobfs_r1,dobfs_twin,dobfs_host_twin >> $out/pull_ab_s${s}_${name}_$tag.jsonl 2>> $out/pull_ab_$tag.err || echo "eval s$s $name FAILED"
    done
  done
  python - "$@" <<PY
import json, sys
for name in sys.argv[1:]:
    for l in open("$out/pull_ab_s${s}_%s_$tag.jsonl" % name):
        r = json.loads(l)
        if "variant" not in r: continue
        print("s$s %-7s %-16s ms=%.4f same=%s" % (name, r["variant"], r["ms"], r["labels_equal_first_variant"]),
              [round(lv["ms"], 4) for lv in r["level_records"] if lv["kind"] == "pull"])
PY
done
