#!/bin/bash
# usage: gpu_r2_scale.sh <N> <outdir> ; headline bench at N GPUs (default settings) + variants
N=$1; O=gpurun_out/${2:-r2s$N}; mkdir -p $O
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --steps 20 --warmup 5"
show() { grep '^{' $1 | tail -1 | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); r=d["roofline"]; print(round(d["ms_per_step"],4),"ms", round(d["value"]/1e9,3),"G/s e2e", round(d["e2e"]["value"]/1e9,3), "vsb", d.get("vs_baseline"), "pulls/step", r.get("rows_pulled_per_step_per_gpu"), "nvl", round(r.get("nvlink_gbs_per_dir_actual",0),1), r.get("nvlink_frac_of_measured_p2p"), "maxif", d.get("max_in_flight"), "nan", d.get("nan_count"), "probe", (d.get("probe") or {}).get("ok"))
except Exception as e: print("FAILED", e)'; }
$T > $O/bench_default.log 2>&1; echo "default: $(show $O/bench_default.log)"
$T --item-dist uniform --baseline-steps 0 > $O/bench_uniform.log 2>&1; echo "uniform: $(show $O/bench_uniform.log)"
$T --variant combine --pull-limit 0 --baseline-steps 0 > $O/bench_combine.log 2>&1; echo "combine(r1): $(show $O/bench_combine.log)"
