#!/bin/bash
O=gpurun_out/r2n256; mkdir -p $O; rm -f $O/*.jsonl
timeout -s KILL 600 python -m pytest tests/test_device_topk.py -m gpu -x -q 2>&1 | tail -4
timeout -s KILL 300 python benchmarks/run_configs.py --configs 6 --out $O/config6.jsonl > $O/c6.log 2>&1; python -c "
import json; r=json.loads(open('$O/config6.jsonl').read().splitlines()[-1]); print({k:v for k,v in r.items() if 'packed' in k or 'error' in k})"
