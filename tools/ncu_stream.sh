#!/bin/bash
# full ncu capture of the streaming advance / filter kernels: tools/ncu_stream.sh <scale> <tag>
out=gpurun_out; mkdir -p $out
scale=${1:-24}; tag=${2:-r02}
python tools/stream_eval.py $scale 1 td_stream_twin > $out/ncu_stream_plain_a.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'StreamExpandKernel' -s 2 -c 2 -f -o $out/prof_stream_expand_$tag \
    python tools/stream_eval.py $scale 1 td_stream_twin > $out/ncu_stream_expand.log 2>&1
echo "expand rc=$?"; tail -2 $out/ncu_stream_expand.log
python tools/stream_eval.py $scale 1 idem_stream_twin > $out/ncu_stream_plain_b.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'filter::StreamKernel' -s 1 -c 1 -f -o $out/prof_stream_filter_$tag \
    python tools/stream_eval.py $scale 1 idem_stream_twin > $out/ncu_stream_filter.log 2>&1
echo "filter rc=$?"; tail -2 $out/ncu_stream_filter.log
