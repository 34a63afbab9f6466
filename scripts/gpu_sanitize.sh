#!/bin/bash
# ONE compute-sanitizer tool (memcheck) on the small kernel tests: tile path incl. the cp.async.bulk rings,
# direct kernel with the limiter, the streaming ring consumer, the sampler, PA kernels
O=gpurun_out/${1:-sanitize}; mkdir -p $O
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_device_mf.py tests/test_device_pa.py -m gpu -q --timeout 800 \
  -k "tiles_variant_equals_oracle or single_walker or bad_ids or negative_sampler or pull_limiter or idle_timeout or padding or pa" > $O/memcheck.log 2>&1
echo "memcheck exit $?" | tee -a $O/memcheck.log
grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds|closed|not supported|refus" $O/memcheck.log | tail -12
