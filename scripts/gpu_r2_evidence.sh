#!/bin/bash
# one-GPU evidence run: tests, headline line, configs, checkpoint timing, ncu of the tile kernel + launch list
O=gpurun_out/${1:-r2j}; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/gpu.txt
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -12
timeout 400 python bench.py --steps 20 --warmup 5 > $O/bench_1.log 2>&1; tail -1 $O/bench_1.log | cut -c1-400
timeout 300 python scripts/bench_checkpoint.py > $O/checkpoint.log 2>&1; tail -1 $O/checkpoint.log
rm -f $O/configs.jsonl
timeout 1200 python benchmarks/run_configs.py --configs 1,2,3,4,6,7,8 --out $O/configs.jsonl > $O/configs.log 2>&1
cut -c1-400 $O/configs.jsonl
P="python bench.py --steps 3 --warmup 2 --baseline-steps 0 --no-probe --no-e2e --variant tiles"
$P > $O/plain_tiles.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_tiles.csv $P > $O/ncu_launch.log 2>&1 \
 && ncu --set full --clock-control none --import-source on -k regex:k_mf_tiles_tma -s 1 -c 1 -o $O/prof_tiles $P > $O/ncu_tiles.log 2>&1
tail -2 $O/ncu_tiles.log
