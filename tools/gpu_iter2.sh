#!/bin/bash
# streaming-advance parity tests (bounded), then a full ncu capture of StreamExpandKernel at the given scale
tag=$1; scale=${2:-25}; variant=${3:-td_stream_twin}
out=gpurun_out; mkdir -p $out
echo skip pytest
python tools/stream_eval.py $scale 1 $variant > $out/ncu_plain_$tag.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'StreamExpandKernel' -s 2 -c 2 -f -o $out/prof_stream_expand_$tag \
    python tools/stream_eval.py $scale 1 $variant > $out/ncu_stream_expand_$tag.log 2>&1
echo "ncu rc=$?"
