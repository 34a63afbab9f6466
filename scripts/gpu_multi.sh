#!/bin/bash
# usage: gpu_multi.sh N   (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR scripts/multi_gpu_check.py > gpurun_out/multi_check_$N.log 2>&1; echo "check exit $?"; grep -E "rank 0\]|MULTI_GPU|p2p probe|Error|error" gpurun_out/multi_check_$N.log | tail -20
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_ours_$N.log 2>&1; echo "bench ours exit $?"; tail -1 gpurun_out/bench_ours_$N.log
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 3 --item-dist uniform --no-e2e > gpurun_out/bench_ours_uniform_$N.log 2>&1; tail -1 gpurun_out/bench_ours_uniform_$N.log
timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 --impl nccl_a2a --batch 1048576 > gpurun_out/bench_a2a_$N.log 2>&1; echo "bench a2a exit $?"; tail -1 gpurun_out/bench_a2a_$N.log
