#!/bin/bash
# last seconds of the round's GPU budget: the remaining users of the row bins (world-1 partitioned PageRank, the
# reference's shared-library programs), the no-alloc entry point after them, smoke
tag=${1:-r02o}
out=gpurun_out; mkdir -p $out
timeout 80 python -m pytest tests/test_partitioned_gpu.py tests/test_shared_lib_gpu.py tests/test_zz_no_alloc.py -m gpu -q -k "pagerank or shared or cc or alloc" --timeout 60 > $out/pytest_rest_$tag.log 2>&1; echo "pytest rc=$?"; tail -4 $out/pytest_rest_$tag.log | cut -c1-300
timeout 30 python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke_$tag.log 2>&1; echo "smoke rc=$?"; tail -1 $out/smoke_$tag.log
