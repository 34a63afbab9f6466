#!/bin/bash
# full GPU suite, then ONE compute-sanitizer tool on the small kernel tests (memcheck)
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
timeout 600 python -m pytest tests -m gpu -q --timeout 200 2>&1 | tail -8
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_device_pa.py tests/test_device_mf.py -m gpu -q --timeout 600 -k "not streaming and not acceptance_and" > gpurun_out/memcheck.log 2>&1; echo "memcheck exit $?"
grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds" gpurun_out/memcheck.log | tail -8
