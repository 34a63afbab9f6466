#!/bin/bash
O=gpurun_out/${1:-r2k}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -8
B="timeout 300 python bench.py --steps 20 --warmup 5 --baseline-steps 0 --no-probe"
for v in "tiles 65536 0" "direct 65536 0"; do
  set -- $v
  $B --variant $1 --pull-limit $2 --tile $3 > $O/bench_$1_$2_$3.log 2>&1
  echo "$v: $(tail -1 $O/bench_$1_$2_$3.log | python -c 'import sys,json
d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4),"ms", round(d["value"]/1e9,3),"G/s e2e", round(d["e2e"]["value"]/1e9,3), "pulls/step", d["roofline"].get("rows_pulled_per_step_per_gpu"))')"
done
rm -f $O/configs.jsonl; timeout 300 python benchmarks/run_configs.py --configs 2 --out $O/configs.jsonl > $O/configs.log 2>&1; cut -c1-300 $O/configs.jsonl
P="python bench.py --steps 3 --warmup 2 --baseline-steps 0 --no-probe --no-e2e --variant tiles"
$P > $O/plain_tiles.log 2>&1 && ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 40 --csv --log-file $O/launches_tiles.csv $P > $O/ncu_launch.log 2>&1
grep -E "k_bucket|k_tile|k_mf" $O/launches_tiles.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | tail -10
