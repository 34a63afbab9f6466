#!/bin/bash
O=gpurun_out/r2umma; mkdir -p $O
timeout -s KILL 120 python scripts/umma_probe.py > $O/umma.log 2>&1; echo "rc=$?"; cat $O/umma.log | tail -14
