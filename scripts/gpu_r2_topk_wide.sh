#!/bin/bash
O=gpurun_out/${1:-r2u}; mkdir -p $O
timeout -s KILL 120 python scripts/topk_wide_check.py config6 > $O/wide_check.log 2>&1; echo "rc=$?" >> $O/wide_check.log
tail -24 $O/wide_check.log
