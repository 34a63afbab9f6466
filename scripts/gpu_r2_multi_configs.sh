#!/bin/bash
# usage: gpu_r2_multi_configs.sh <N> <outdir>: correctness check, symmetric probe, configs 2/3/4 at N GPUs
N=$1; O=gpurun_out/${2:-r2m$N}; mkdir -p $O
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $R --master-port 29641 scripts/multi_gpu_check.py > $O/multi_check_$N.log 2>&1; grep -E "MULTI_GPU_CHECK|False" $O/multi_check_$N.log | tail -5
timeout 300 $R --master-port 29642 scripts/p2p_symmetric_probe.py 2>&1 | grep "GB/s" | tee $O/p2p_symmetric_$N.log
rm -f $O/configs.jsonl
timeout 900 $R --master-port 29643 benchmarks/run_configs.py --configs 2,3,4 --out $O/configs.jsonl > $O/configs.log 2>&1
cut -c1-700 $O/configs.jsonl
FPS_CONFIG2_LR=0.01 timeout 300 $R --master-port 29644 benchmarks/run_configs.py --configs 2 --out $O/configs_lr01.jsonl > $O/configs_lr01.log 2>&1
cut -c1-500 $O/configs_lr01.jsonl
