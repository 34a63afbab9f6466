#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
for b in 0 6 8 12 24 64 200; do
  timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --users 20000000 --blocks-per-sm $b 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('bps',$b,'ms',round(d['ms_per_step'],3))"
done
for u in 1 4; do
  timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --users 20000000 --blocks-per-sm 200 --unroll $u 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('bps 200 unroll',$u,'ms',round(d['ms_per_step'],3))"
done
