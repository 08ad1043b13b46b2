#!/bin/bash
# full ncu capture of filter::Kernel on the big level of an idempotent BFS
out=gpurun_out; mkdir -p $out
cat > /tmp/filter_probe.py <<'P'
import sys, os
sys.path.insert(0, os.getcwd())
import gunrockinst_b200 as gr
g = gr.Graph.rmat(int(sys.argv[1]))
bfs = gr.Bfs(g, idempotence=True)
for i in range(2):
    bfs.run(0, want_labels=False)
P
python /tmp/filter_probe.py ${1:-22} > $out/filter_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'filter::Kernel' -s 7 -c 1 -f -o $out/prof_filter_${2:-r01} python /tmp/filter_probe.py ${1:-22} > $out/ncu_filter.log 2>&1
echo "rc=$?"; tail -2 $out/ncu_filter.log
