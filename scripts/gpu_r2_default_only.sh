#!/bin/bash
N=$1; O=gpurun_out/${2:-r2z$N}; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29661 bench.py --gpus $N --steps 50 --warmup 10 > $O/bench_default.log 2>&1
grep '^{' $O/bench_default.log | tail -1 | python -c 'import sys,json
d=json.loads(sys.stdin.read()); r=d["roofline"]; print(d["n_gpus"], round(d["ms_per_step"],4),"ms", round(d["value"]/1e9,3),"G/s e2e", round(d["e2e"]["value"]/1e9,3), "vsb", round(d["vs_baseline"],2), "pulls/step", r["rows_pulled_per_step_per_gpu"], "nvl", round(r["nvlink_gbs_per_dir_actual"],1), round(r["nvlink_frac_of_symmetric_probe"],3), "limit", d["config"]["pull_limit"], "maxif", d["max_in_flight"], "probe", d["probe"]["ok"], d["training_valid"], d["clocks"])'
