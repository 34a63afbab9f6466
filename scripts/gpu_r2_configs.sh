#!/bin/bash
O=gpurun_out/${1:-r2g}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -12
rm -f $O/configs.jsonl
timeout 900 python benchmarks/run_configs.py --configs ${2:-1,2,3,4} --out $O/configs.jsonl > $O/configs.log 2>&1
cut -c1-900 $O/configs.jsonl
FPS_CONFIG2_LR=0.01 timeout 300 python benchmarks/run_configs.py --configs 2 --out $O/configs_lr01.jsonl > $O/configs_lr01.log 2>&1
cut -c1-700 $O/configs_lr01.jsonl
