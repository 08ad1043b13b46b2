#!/bin/bash
out=gpurun_out; mkdir -p $out
for d in 0 16 32 64 96; do
GRB_STREAM_DEBUG=$d timeout 60 python tools/stream_eval.py 25 2 td_stream_twin > $out/dbg_$d.jsonl 2>/dev/null
python - <<PY
import json
for l in open("$out/dbg_$d.jsonl"):
    if not l.startswith("{"): continue
    r=json.loads(l)
    if "variant" not in r: continue
    print("debug=$d ms=%.3f"%(r["ms"]), [ (lv["kind"],lv["edges"],lv["discovered"],round(lv["ms"],3)) for lv in r["level_records"] if lv["ms"]>0.05][:4])
PY
done
