#!/bin/bash
# round 2, first GPU call: GPU tests, then the tile variant against the direct kernel on one GPU
O=gpurun_out/r2a; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total --format=csv > $O/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -5 $O/pytest.log
B="timeout 300 python bench.py --steps 20 --warmup 5"
$B > $O/bench_default.log 2>&1; tail -1 $O/bench_default.log | cut -c1-600
for v in "direct 0 0" "tiles 0 0" "tiles 0 512" "tiles 0 2048" "tiles 65536 0" "tiles 1600 0" "direct 65536 0" "direct 1600 0" "combine 0 0"; do
  set -- $v
  $B --variant $1 --pull-limit $2 --tile $3 --baseline-steps 0 --no-probe > $O/bench_$1_$2_$3.log 2>&1
  echo "$v: $(tail -1 $O/bench_$1_$2_$3.log | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4),"ms", round(d["value"]/1e9,3),"G/s e2e", round(d["e2e"]["value"]/1e9,3), "pulls/step", d["roofline"].get("rows_pulled_per_step_per_gpu"), "hbm", round(d["roofline"].get("hbm_frac_of_measured_copy_peak",0),3), "maxif", d.get("max_in_flight"), "nan", d.get("nan_count"))
except Exception as e: print("FAILED", e)')"
done
