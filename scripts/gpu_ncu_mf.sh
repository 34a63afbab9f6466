#!/bin/bash
# one full ncu capture of a fused MF kernel (run only after the plain command passed)
#   bash scripts/gpu_ncu_mf.sh direct k_mf_sgd      |  bash scripts/gpu_ncu_mf.sh combine k_mf_sgd_combine
mkdir -p gpurun_out
VARIANT=${1:-direct}
KERNEL=${2:-k_mf_sgd}
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --users 20000000 --items 2000000 --variant $VARIANT"
timeout 600 $CMD > gpurun_out/plain_$VARIANT.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KERNEL -s 3 -c 1 -f -o gpurun_out/prof_mf_$VARIANT $CMD > gpurun_out/ncu_full_$VARIANT.log 2>&1
tail -n 2 gpurun_out/plain_$VARIANT.log; tail -n 3 gpurun_out/ncu_full_$VARIANT.log
