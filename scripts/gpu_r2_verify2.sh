#!/bin/bash
# 2 GPUs: whole GPU test suite (incl. the 2-rank check), default bench, config 4
O=gpurun_out/${1:-r2v2}; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -8
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 > $O/bench2.log 2>&1; grep '^{' $O/bench2.log | tail -1 | cut -c1-420
rm -f $O/configs.jsonl; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 benchmarks/run_configs.py --configs 3,4 --out $O/configs.jsonl > $O/configs.log 2>&1; cut -c1-100 $O/configs.jsonl; python -c "
import json
for l in open('$O/configs.jsonl'):
    r=json.loads(l); print(r.get('config'), {k:r.get(k) for k in ('ms_per_step','without_limiter_ms_per_step','examples_per_sec','error')})"
