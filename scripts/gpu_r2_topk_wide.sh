#!/bin/bash
O=gpurun_out/${1:-r2u}; mkdir -p $O
timeout -s KILL 150 python scripts/topk_wide_check.py config6 > $O/wide_check.log 2>&1; echo "rc=$?" >> $O/wide_check.log
tail -8 $O/wide_check.log
FPS_TOPK_FILL=0 timeout -s KILL 150 python scripts/topk_wide_check.py config6 > $O/wide_check_nofill.log 2>&1; echo "rc=$?" >> $O/wide_check_nofill.log
tail -3 $O/wide_check_nofill.log
timeout -s KILL 300 python -m pytest tests/test_device_topk.py tests/test_zz_no_alloc.py tests/test_device_jobs.py -m gpu -q > $O/pytest_topk.log 2>&1; echo "pytest rc=$?" >> $O/pytest_topk.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest_topk.log | tail -6
