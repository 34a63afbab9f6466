#!/bin/bash
# 2-GPU call: full GPU test suite (incl. the 2-rank check), then the headline bench
O=gpurun_out/${1:-r2f}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -12
T="timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5"
for v in "tiles 65536 0" "tiles 0 0" "tiles 0 2048"; do
  set -- $v
  $T --variant $1 --pull-limit $2 --tile $3 --baseline-steps 0 > $O/bench2_$1_$2_$3.log 2>&1
  echo "$v: $(grep '^{' $O/bench2_$1_$2_$3.log | tail -1 | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); r=d["roofline"]; print(round(d["ms_per_step"],4),"ms", round(d["value"]/1e9,3),"G/s e2e", round(d["e2e"]["value"]/1e9,3), "pulls/step", r.get("rows_pulled_per_step_per_gpu"), "nvl", round(r.get("nvlink_gbs_per_dir_actual",0),1), r.get("nvlink_frac_of_measured_p2p"), "maxif", d.get("max_in_flight"), "nan", d.get("nan_count"), "probe", (d.get("probe") or {}).get("ok"))
except Exception as e: print("FAILED", e)')"
done
