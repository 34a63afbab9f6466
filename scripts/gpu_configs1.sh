#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
timeout 600 python benchmarks/run_configs.py --configs 1,2,3,4,6 2>&1 | grep -E "^\{|rror" | cut -c1-900
python scripts/topk_once.py > gpurun_out/plain_topk.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_lemp_topk -s 2 -c 1 -o gpurun_out/prof_topk python scripts/topk_once.py > gpurun_out/ncu_topk.log 2>&1
tail -2 gpurun_out/ncu_topk.log
