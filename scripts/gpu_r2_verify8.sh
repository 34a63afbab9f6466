#!/bin/bash
# 8 GPUs: config 3 after the binary PA kernel change
O=gpurun_out/${1:-r2v8}; mkdir -p $O
rm -f $O/configs.jsonl; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29612 benchmarks/run_configs.py --configs 3 --out $O/configs.jsonl > $O/configs.log 2>&1; python -c "
import json
for l in open('$O/configs.jsonl'):
    r=json.loads(l); print(r.get('config'), {k:r.get(k) for k in ('ms_per_step','examples_per_sec','table','without_limiter','error')})"
