#!/bin/bash
# compute-sanitizer evidence: plain run first, then memcheck, racecheck and synccheck of the same probe
tag=${1:-r02}; scale=${2:-12}
out=gpurun_out; mkdir -p $out
timeout 300 python tools/sanitize_probe.py $scale > $out/sanitizer_${tag}_plain.log 2>&1; echo "plain rc=$?"; tail -2 $out/sanitizer_${tag}_plain.log
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 40 --log-file $out/sanitizer_${tag}_$tool.log python tools/sanitize_probe.py $scale > $out/sanitizer_${tag}_${tool}_stdout.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|ALL OK|FAILURES" $out/sanitizer_${tag}_$tool.log $out/sanitizer_${tag}_${tool}_stdout.log | tail -4
done
