#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
r() { timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --users 20000000 "$@" 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*','ms',round(d['ms_per_step'],3))"; }
r --unroll 1 --blocks-per-sm 6
r --unroll 1 --blocks-per-sm 12
r --unroll 1 --blocks-per-sm 48
r --unroll 1 --blocks-per-sm 200
r --unroll 1 --blocks-per-sm 8 --min-blocks 8
r --unroll 1 --blocks-per-sm 16 --min-blocks 8
r --unroll 1 --blocks-per-sm 200 --min-blocks 8
r --unroll 1 --blocks-per-sm 200 --item-dist uniform
r --unroll 1 --blocks-per-sm 200 --min-blocks 8 --item-dist uniform
