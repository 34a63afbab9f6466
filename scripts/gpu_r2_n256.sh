#!/bin/bash
for g in 1 2 4; do echo "64-item kernel, groups $g:"; FPS_TOPK_WIDE=0 FPS_TOPK_GROUPS=$g timeout -s KILL 300 python scripts/topk_rank_sweep.py 2>&1 | grep rank; done
