#!/bin/bash
# one full ncu capture of a PA kernel (config 3 = k_pa_binary, config 4 = k_pa_multi), after the plain run passed
mkdir -p gpurun_out
CFG=${1:-4}
KERNEL=${2:-k_pa_multi}
CMD="python benchmarks/run_configs.py --configs $CFG"
timeout 300 $CMD > gpurun_out/pa_plain_$CFG.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$KERNEL -s 5 -c 1 -f -o gpurun_out/prof_pa_$CFG $CMD > gpurun_out/ncu_pa_$CFG.log 2>&1
grep config gpurun_out/pa_plain_$CFG.log | cut -c1-400; tail -n 2 gpurun_out/ncu_pa_$CFG.log
