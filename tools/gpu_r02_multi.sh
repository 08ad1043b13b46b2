#!/bin/bash
# round-2 session on one 8-GPU box: partitioned parity at world 2 / 4 / 8 (all four primitives; CC at 4 / 8 for the first
# time; BFS through both level loops), the scale-27 BFS scaling points the way the driver launches them (torchrun, N = 1, 2,
# 4 side by side on disjoint devices, then 8), the N = 8 point with the search as one cooperative kernel, and the
# partitioned CC table at scale 25.
tag=${1:-r02}
out=gpurun_out; mkdir -p $out
export GRB_PEER_TIMEOUT_MS=8000
nvidia-smi -L | head -8 > $out/smi_multi_$tag.log
timeout 1500 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q > $out/pytest_partitioned_8gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -4 $out/pytest_partitioned_8gpu_$tag.log
run_n() {   # run_n <n> <devices> <port> <extra args...>
  n=$1; devs=$2; port=$3; shift 3
  CUDA_VISIBLE_DEVICES=$devs timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $n --steps 64 --warmup 3 "$@"
}
run_n 1 0 29601 > $out/scale_${tag}_n1.json 2> $out/scale_${tag}_n1.err &
run_n 2 1,2 29602 > $out/scale_${tag}_n2.json 2> $out/scale_${tag}_n2.err &
run_n 4 3,4,5,6 29603 > $out/scale_${tag}_n4.json 2> $out/scale_${tag}_n4.err &
wait
run_n 8 0,1,2,3,4,5,6,7 29604 > $out/scale_${tag}_n8.json 2> $out/scale_${tag}_n8.err; echo "n8 rc=$?"
run_n 8 0,1,2,3,4,5,6,7 29605 --mg-persistent --no-single-ref > $out/scale_${tag}_n8_persistent.json 2> $out/scale_${tag}_n8_persistent.err; echo "n8 persistent rc=$?"
run_n 2 0,1 29606 --mg-persistent --no-single-ref > $out/scale_${tag}_n2_persistent.json 2> $out/scale_${tag}_n2_persistent.err &
run_n 4 2,3,4,5 29607 --mg-persistent --no-single-ref > $out/scale_${tag}_n4_persistent.json 2> $out/scale_${tag}_n4_persistent.err &
wait
python - <<PY
import json
for f in ["n1","n2","n4","n8","n8_persistent","n2_persistent","n4_persistent"]:
    try:
        r=json.loads(open("$out/scale_${tag}_%s.json"%f).read().strip().splitlines()[-1])
        s=r.get("single_device_same_graph") or {}
        print(f, "gteps=%.0f ms=%.3f launches/search/gpu=%.1f parity=%s single=%s speedup=%s exch=%.3f comp=%.3f e2e=%.0f"%(r["value"], r["ms_per_step"], r["gpu_launches"]/r["steps"]/r["n_gpus"], (r.get("parity") or {}).get("ok"), s.get("gteps"), r.get("speedup_vs_single_device"), r["exchange"]["ms_per_search"], r["exchange"]["compute_ms_per_search"], r["e2e"]["value"]))
    except Exception as e:
        print(f, "failed", e)
PY
for n in 1 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2962$n tools/mg_cc_bench.py 25 3 > $out/mgcc_${tag}_n${n}_s25.json 2> $out/mgcc_${tag}_n${n}_s25.err
  echo "mgcc n=$n rc=$?"; cut -c1-400 $out/mgcc_${tag}_n${n}_s25.json
done
