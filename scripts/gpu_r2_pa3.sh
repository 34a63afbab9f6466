#!/bin/bash
O=gpurun_out/r2pa3; mkdir -p $O; rm -f $O/config3.jsonl
timeout -s KILL 900 python -m pytest tests/test_device_pa.py tests/test_device_dispatch.py -m gpu -x -q 2>&1 | tail -2
timeout -s KILL 300 python benchmarks/run_configs.py --configs 3 --out $O/config3.jsonl > $O/c3.log 2>&1; python -c "
import json
for l in open('$O/config3.jsonl'):
    r=json.loads(l); print(r.get('ms_per_step'), r.get('examples_per_sec'), r.get('without_limiter'), r.get('error'))"
