#!/bin/bash
O=gpurun_out/r2b; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -4 $O/pytest.log
B="timeout 300 python bench.py --steps 20 --warmup 5 --baseline-steps 0 --no-probe"
for v in "direct 0 0" "direct 65536 0" "direct 1600 0" "tiles 0 0" "tiles 65536 0" "tiles 1600 0"; do
  set -- $v
  $B --variant $1 --pull-limit $2 --tile $3 > $O/bench_$1_$2_$3.log 2>&1
  echo "$v: $(tail -1 $O/bench_$1_$2_$3.log | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4),"ms", round(d["value"]/1e9,3),"G/s e2e", round(d["e2e"]["value"]/1e9,3), "pulls/step", d["roofline"].get("rows_pulled_per_step_per_gpu"), "hbm", round(d["roofline"].get("hbm_frac_of_measured_copy_peak",0),3), "maxif", d.get("max_in_flight"), "nan", d.get("nan_count"))
except Exception as e: print("FAILED", e)')"
done
P="python bench.py --steps 2 --warmup 1 --baseline-steps 0 --no-probe --no-e2e --variant tiles --pull-limit 0"
$P > $O/plain_tiles.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_mf_tiles_tma -s 1 -c 1 -o $O/prof_tiles $P > $O/ncu_tiles.log 2>&1
tail -2 $O/ncu_tiles.log
