#!/bin/bash
# launch list (one ncu pass, gpu__time_duration only) of the row-binned passes at scale 25: binning, hook, flatten, sweeps
out=gpurun_out; mkdir -p $out
timeout 90 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled \
    -k regex:'BinKernel|CcRowHook|CcFlatten|PrRowSweep|PrBigRow|MgPrStart|ExchangeCounters' --csv \
    --log-file $out/launches_rows_r02p.csv python tools/rows_ab.py 25 4 1 > $out/ncu_launches_rows_r02p.log 2>&1
echo "rc=$?"; wc -l $out/launches_rows_r02p.csv; tail -2 $out/ncu_launches_rows_r02p.log | cut -c1-300
