#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 400 $TR scripts/mf_2gpu_diag.py > gpurun_out/mf_diag.log 2>&1; echo "diag exit $?"; grep -E "ms|rror" gpurun_out/mf_diag.log | tail -40
DIAG_USERS=2000000 timeout 300 $TR scripts/mf_2gpu_diag.py > gpurun_out/mf_diag_small.log 2>&1; echo "diag small exit $?"; grep -E "ms/step|rror" gpurun_out/mf_diag_small.log | tail -10
timeout 300 python -m pytest tests/test_device_jobs.py -m gpu -q --timeout 120 2>&1 | tail -15
