#!/bin/bash
out=gpurun_out; mkdir -p $out
for i in 1 2 3; do timeout 100 python tools/idem_diff.py 22 | grep -E "mismatch"; done
timeout 100 python tools/idem_diff.py 25 | grep -E "mismatch"
bash tools/gpu_iter.sh r02u td_stream_twin,td_stream,idem_r1,idem_stream,idem_stream_twin 25 22
