#!/bin/bash
# launch list + one full ncu capture of the fused MF kernel (run only after the plain command passed)
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --users 20000000 --items 2000000"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_mf.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_mf_sgd -s 3 -c 1 -o gpurun_out/prof_mf $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/
