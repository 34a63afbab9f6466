#!/bin/bash
# usage: gpu_r2_config2.sh <N> <outdir> [limits...]: GPU tests, then config 2 at N GPUs under several pull limits / learning rates
N=$1; O=gpurun_out/${2:-r2c2_$N}; mkdir -p $O; shift; shift
if [ "$N" -le 2 ]; then
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -8
fi
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
rm -f $O/config2.jsonl
for spec in "$@"; do
  lim=${spec%%:*}; lr=${spec##*:}
  FPS_CONFIG2_PULL_LIMIT=$lim FPS_CONFIG2_LR=$lr timeout 300 $R --master-port 29651 benchmarks/run_configs.py --configs 2 --out $O/config2.jsonl > $O/config2_$lim_$lr.log 2>&1
done
python - <<PY
import json
for l in open("$O/config2.jsonl"):
    d=json.loads(l)
    print({k:d.get(k) for k in ("n_gpus","pull_limit","learning_rate","ms_per_step","ratings_per_sec","nan_count","mse","max_in_flight","rows_pulled_per_record","error")})
PY
