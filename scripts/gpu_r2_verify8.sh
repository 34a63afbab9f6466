#!/bin/bash
# 8 GPUs: configs 3 and 4 after the PA kernel changes, the 8-rank correctness check
O=gpurun_out/${1:-r2v8}; mkdir -p $O
rm -f $O/configs.jsonl; timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29612 benchmarks/run_configs.py --configs 3,4 --out $O/configs.jsonl > $O/configs.log 2>&1; python -c "
import json
for l in open('$O/configs.jsonl'):
    r=json.loads(l); print(r.get('config'), {k:r.get(k) for k in ('ms_per_step','without_limiter_ms_per_step','examples_per_sec','table','error')})"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29613 scripts/multi_gpu_check.py > $O/multi_check_8.log 2>&1; tail -3 $O/multi_check_8.log | cut -c1-200
