#!/bin/bash
# full ncu capture of the pull sweep (host-driven loop so that every level is its own kernel)
out=gpurun_out; mkdir -p $out
cat > /tmp/pull_probe.py <<'P'
import sys, os
sys.path.insert(0, os.getcwd())
import gunrockinst_b200 as gr
g = gr.Graph.rmat(int(sys.argv[1]))
bfs = gr.Bfs(g, direction_optimized=True, host_loop=True)
for i in range(3):
    bfs.run(0, want_labels=False)
P
python /tmp/pull_probe.py ${1:-22} > $out/pull_plain.log 2>&1 || exit 1
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'pull::Kernel' -s 3 -c 3 -f -o $out/prof_pull_${2:-r01c} python /tmp/pull_probe.py ${1:-22} > $out/ncu_pull.log 2>&1
echo "rc=$?"; tail -3 $out/ncu_pull.log
