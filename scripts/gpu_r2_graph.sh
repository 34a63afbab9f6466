#!/bin/bash
O=gpurun_out/r2graph; mkdir -p $O; rm -f $O/*.jsonl
timeout -s KILL 900 python -m pytest tests/test_device_jobs.py -m gpu -x -q 2>&1 | tail -3
