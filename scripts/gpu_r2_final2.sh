#!/bin/bash
O=gpurun_out/${1:-r2x}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -8
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_2.log 2>&1
grep '^{' $O/bench_2.log | tail -1 | python -c 'import sys,json
d=json.loads(sys.stdin.read()); r=d["roofline"]; print(round(d["ms_per_step"],4),"ms", round(d["value"]/1e9,3),"G/s e2e", round(d["e2e"]["value"]/1e9,3), "vsb", round(d["vs_baseline"],2), "pulls/step", r["rows_pulled_per_step_per_gpu"], "nvl", round(r["nvlink_gbs_per_dir_actual"],1), round(r["nvlink_frac_of_symmetric_probe"],3), "limit", d["config"]["pull_limit"], "maxif", d["max_in_flight"], "probe", d["probe"]["ok"], d["training_valid"])'
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_1.log 2>&1; tail -1 $O/bench_1.log | cut -c1-330
