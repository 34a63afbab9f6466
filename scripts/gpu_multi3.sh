#!/bin/bash
# usage: gpu_multi3.sh N   (run under gpurun --gpus N): correctness check + default bench + pytest multi-gpu
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR scripts/multi_gpu_check.py > gpurun_out/multi_check_$N.log 2>&1; echo "check exit $?"; grep -E "MULTI_GPU|Error|error" gpurun_out/multi_check_$N.log | tail -5
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_ours_$N.log 2>&1; echo "bench ours exit $?"; tail -n 1 gpurun_out/bench_ours_$N.log | cut -c1-900
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -x 2>&1 | tail -n 3
