#!/bin/bash
# microbenchmark of the id-stream + visited-probe shapes, then full ncu captures of the streaming kernels at scale 25
out=gpurun_out; mkdir -p $out
timeout 300 ./build/probe_micro 25 28 > $out/probe_micro_s25.json 2> $out/probe_micro_s25.err; echo "micro rc=$?"
bash tools/ncu_stream.sh 25 r02b
python tools/stream_eval.py 25 1 td_r1_twin > $out/ncu_r1_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'lb::ExpandKernel' -s 2 -c 2 -f -o $out/prof_r1_expand_r02b \
    python tools/stream_eval.py 25 1 td_r1_twin > $out/ncu_r1_expand.log 2>&1
echo "r1 expand rc=$?"
