#!/bin/bash
# one GPU: whole GPU test suite, smoke, default bench, configs 3 and 4
O=gpurun_out/${1:-r2v}; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -8
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py > $O/bench_1_defaults.log 2>&1; tail -1 $O/bench_1_defaults.log | cut -c1-400
rm -f $O/configs.jsonl; timeout 600 python benchmarks/run_configs.py --configs 3,4 --out $O/configs.jsonl > $O/configs.log 2>&1; cut -c1-600 $O/configs.jsonl
