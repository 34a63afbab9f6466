#!/bin/bash
# one full ncu capture of the packed tcgen05 top-K kernel (only after the plain command passed)
mkdir -p gpurun_out
MODE=${1:-config6}
CMD="python scripts/topk_once.py $MODE"
timeout 300 $CMD > gpurun_out/topk_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lemp_topk_packed -s 2 -c 1 -f -o gpurun_out/prof_topk_packed_$MODE $CMD > gpurun_out/ncu_topk.log 2>&1
cat gpurun_out/topk_plain.log; tail -n 3 gpurun_out/ncu_topk.log
