#!/bin/bash
out=gpurun_out; mkdir -p $out
export GRB_PEER_TIMEOUT_MS=5000
timeout 600 python -m pytest tests/test_partitioned_gpu.py -m gpu -x -q -k "world1 or multi_device_matches_oracle and not sssp and not pagerank and not cc" > $out/pytest_g.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest_g.log
