#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
timeout 300 python -m pytest tests/test_device_mf.py -m gpu -q --timeout 150 2>&1 | tail -4
r() { timeout 300 python bench.py --steps 20 --warmup 3 "$@" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*','ms',round(d['ms_per_step'],3),'G/s',round(d['value']/1e9,3),'e2e',d['e2e'] and round(d['e2e']['value']/1e9,3),'mse',round(d['mse'],4))"; }
r
r --item-dist uniform
r --variant combine
r --variant combine --item-dist uniform
r --pull-limit 1600
