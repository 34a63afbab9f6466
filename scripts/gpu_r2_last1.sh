#!/bin/bash
O=gpurun_out/${1:-r2y}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -8
timeout 600 python bench.py --impl cpu --steps 5 --warmup 2 > $O/bench_cpu.log 2>&1; tail -1 $O/bench_cpu.log | cut -c1-500
rm -f $O/configs.jsonl; timeout 600 python benchmarks/run_configs.py --configs 2,6 --out $O/configs.jsonl > $O/configs.log 2>&1; cut -c1-420 $O/configs.jsonl
timeout 300 python bench.py > $O/bench_1_defaults.log 2>&1; tail -1 $O/bench_1_defaults.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
