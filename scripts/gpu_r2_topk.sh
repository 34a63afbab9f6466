#!/bin/bash
O=gpurun_out/${1:-r2t}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -8
CMD="python scripts/topk_once.py config6"
timeout 300 $CMD > $O/topk_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lemp_topk_packed -s 2 -c 1 -f -o $O/prof_topk_packed $CMD > $O/ncu_topk.log 2>&1
cat $O/topk_plain.log; tail -n 2 $O/ncu_topk.log
