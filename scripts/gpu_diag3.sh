#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1
timeout 200 python scripts/mf_peer_single_process.py 2>&1 | tail -5
ONLY=all_remote timeout 200 python scripts/mf_peer_single_process.py > gpurun_out/plain_peer.log 2>&1 && \
ONLY=all_remote timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_mf_sgd -s 3 -c 1 -o gpurun_out/prof_mf_remote python scripts/mf_peer_single_process.py > gpurun_out/ncu_remote.log 2>&1
tail -3 gpurun_out/ncu_remote.log
