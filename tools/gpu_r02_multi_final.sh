#!/bin/bash
# closing 8-GPU session of round 2: BFS parity at world 2 / 4 / 8 through every level loop (host, one kernel, host + TAIL) and
# on directed slices, then the scale-27 scaling points as the driver launches them (with parity, the single-device search of
# the same graph, e2e and e2e_from_host)
tag=${1:-r02z}
out=gpurun_out; mkdir -p $out
export GRB_PEER_TIMEOUT_MS=8000
timeout 900 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q -k "test_multi_device_matches_oracle or test_cc_multi" > $out/pytest_partitioned_8gpu_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_partitioned_8gpu_$tag.log
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
python - <<PY
import json
for f in ["n1","n2","n4","n8"]:
    try:
        r=json.loads(open("$out/scale_${tag}_%s.json"%f).read().strip().splitlines()[-1])
        s=r.get("single_device_same_graph") or {}
        print(f, "gteps=%.0f ms=%.3f parity=%s single=%s speedup=%s exch=%.3f comp=%.3f e2e=%.0f from_host=%s"%(r["value"], r["ms_per_step"], (r.get("parity") or {}).get("ok"), s.get("gteps"), r.get("speedup_vs_single_device"), r["exchange"]["ms_per_search"], r["exchange"]["compute_ms_per_search"], r["e2e"]["value"], (r.get("e2e_from_host") or {}).get("value")))
    except Exception as e:
        print(f, "failed", e)
PY
