#!/bin/bash
# usage: gpu_multi2.sh N : direct vs combine, skewed vs uniform
N=${1:-8}
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
r() { tag=$1; shift; timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 3 "$@" > gpurun_out/bench_${tag}_$N.log 2>&1; tail -1 gpurun_out/bench_${tag}_$N.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$tag N=$N','ms',round(d['ms_per_step'],3),'G/s',round(d['value']/1e9,3),'e2e',d['e2e'] and round(d['e2e']['value']/1e9,3),'mse',round(d['mse'],4),'nvlink',d['roofline']['nvlink_frac_of_770'])" || tail -5 gpurun_out/bench_${tag}_$N.log; }
r direct_ml
r direct_uni --item-dist uniform --no-e2e
r combine_ml --variant combine
r combine_uni --variant combine --item-dist uniform --no-e2e
