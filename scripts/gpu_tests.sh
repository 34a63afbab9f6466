#!/bin/bash
# GPU test sweep, one file at a time under its own timeout so one hang costs one file only
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
for f in tests/test_device_mf.py tests/test_device_pa.py tests/test_device_topk.py; do
  n=$(basename $f .py)
  timeout 420 python -m pytest $f -m gpu -q --timeout 150 > gpurun_out/$n.log 2>&1
  echo "== $n exit $?"; tail -25 gpurun_out/$n.log
done
