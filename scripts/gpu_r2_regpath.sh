#!/bin/bash
O=gpurun_out/${1:-r2n}; mkdir -p $O
B="timeout 300 python bench.py --steps 20 --warmup 5 --baseline-steps 0 --no-probe"
for v in "tiles 0 0" "tiles 1 0" "tiles 1 512" "tiles 1 2048"; do
  set -- $v
  $B --variant $1 --unroll $2 --tile $3 > $O/bench_$1_$2_$3.log 2>&1
  echo "$v: $(tail -1 $O/bench_$1_$2_$3.log | python -c 'import sys,json
d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"],4),"ms", round(d["value"]/1e9,3),"G/s e2e", round(d["e2e"]["value"]/1e9,3), "pulls/step", d["roofline"].get("rows_pulled_per_step_per_gpu"))')"
done
