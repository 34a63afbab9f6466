#!/bin/bash
O=gpurun_out/r2filters; mkdir -p $O
timeout -s KILL 600 python -m pytest tests/test_device_jobs.py -m gpu -x -q > $O/pytest.log 2>&1; echo "rc=$?"; tail -25 $O/pytest.log | cut -c1-220
