#!/bin/bash
# closing 1-GPU session of round 2: whole GPU suite, smoke, both bench arms, launch list of the bench command, full ncu
# captures of the dominant kernel of the bench step (pull sweep, scale 22) and of the streaming advance (scale 25, twin)
tag=${1:-r02z}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_gpu_$tag.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke_$tag.log 2>&1; echo "smoke rc=$?"; tail -1 $out/smoke_$tag.log
timeout 1500 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?"; cut -c1-600 $out/bench_$tag.json; tail -3 $out/bench_$tag.err
timeout 900 python bench.py --impl reference --steps 32 --warmup 3 > $out/bench_reference_$tag.json 2> $out/bench_reference_$tag.err; echo "ref rc=$?"; cut -c1-400 $out/bench_reference_$tag.json
small="python bench.py --steps 2 --warmup 1 --no-cpu --e2e-steps 1 --no-configs"
$small > $out/plain_small.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 600 --csv --log-file $out/launches_bench_steps2_$tag.csv $small > $out/ncu_launches.log 2>&1
echo "launch list rc=$?"
bash tools/ncu_pull.sh 22 $tag
python tools/stream_eval.py 25 1 td_stream_twin > $out/ncu_plain_$tag.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'StreamExpandKernel|HarvestKernel' -s 4 -c 4 -f -o $out/prof_stream_expand_$tag \
    python tools/stream_eval.py 25 1 td_stream_twin > $out/ncu_stream_expand_$tag.log 2>&1
echo "stream capture rc=$?"
python tools/stream_eval.py 25 1 idem_stream_twin > $out/ncu_plain_f_$tag.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:'StreamExpandKernel' -s 1 -c 1 -f -o $out/prof_stream_filter_$tag \
    python tools/stream_eval.py 25 1 idem_stream_twin > $out/ncu_stream_filter_$tag.log 2>&1
echo "filter capture rc=$?"
timeout 300 python tools/stream_eval.py 25 3 td_r1,td_r1_twin,td_stream_twin,dobfs_r1,dobfs_twin,idem_r1,idem_stream_twin > $out/stream_eval_s25_$tag.jsonl 2>/dev/null
timeout 300 python tools/stream_eval.py 22 3 td_r1,td_r1_twin,td_stream_twin,dobfs_r1,dobfs_twin,idem_r1,idem_stream_twin > $out/stream_eval_s22_$tag.jsonl 2>/dev/null
echo "evals done"
