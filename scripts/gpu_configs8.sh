#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR scripts/multi_gpu_check.py > gpurun_out/multi_check_$N.log 2>&1; echo "check exit $?"; grep -E "MULTI_GPU|rror" gpurun_out/multi_check_$N.log | tail -3
timeout 600 $TR benchmarks/run_configs.py --configs 2,3,4 --out gpurun_out/configs_$N.jsonl 2>&1 | grep -E "^\{|rror" | cut -c1-700
timeout 300 python -m pytest tests/test_multi_gpu.py -m gpu -q --timeout 500 2>&1 | tail -3
