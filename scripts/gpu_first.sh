#!/bin/bash
# first GPU contact: tests, smoke, short bench
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -40 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -5 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --users 20000000 --items 2000000 > gpurun_out/bench_small.log 2>&1; echo "bench small exit $?"; tail -3 gpurun_out/bench_small.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo "bench full exit $?"; tail -3 gpurun_out/bench_full.log
timeout 600 python bench.py --steps 20 --warmup 3 --item-dist uniform --no-e2e > gpurun_out/bench_uniform.log 2>&1; tail -1 gpurun_out/bench_uniform.log
timeout 600 python bench.py --steps 10 --warmup 3 --impl nccl_a2a --batch 1048576 > gpurun_out/bench_a2a.log 2>&1; tail -2 gpurun_out/bench_a2a.log
