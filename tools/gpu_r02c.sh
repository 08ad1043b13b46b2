#!/bin/bash
# full GPU parity suite + launch list of one top-down scale-25 search on the twin graph (per-kernel durations)
tag=${1:-r02c}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -4 $out/pytest_gpu_$tag.log
python tools/stream_eval.py 25 1 td_stream_twin > $out/plain_$tag.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 400 --csv --log-file $out/launches_td25_$tag.csv python tools/stream_eval.py 25 1 td_stream_twin > $out/ncu_l_$tag.log 2>&1
echo "launch list rc=$?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("$out/launches_td25_$tag.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value")
for r in rows[-40:]:
    print(r[ki][:90], r[vi])
PY
