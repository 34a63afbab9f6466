#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
(cd ab_old && python build_native.py > /dev/null 2>&1)
L=flink-parameter-server-1_b200/ops/_lib/libfps_b200.so
run() { timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --users 20000000 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1','ms',round(d['ms_per_step'],3))"; }
run "new py + new lib"
cp $L /tmp/new.so; cp ab_old/$L $L
run "new py + old lib"
cp /tmp/new.so $L
(cd ab_old && cp $L /tmp/old.so && cp /tmp/new.so $L && timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --users 20000000 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('old py + new lib','ms',round(d['ms_per_step'],3))"; cp /tmp/old.so $L; timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --users 20000000 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('old py + old lib','ms',round(d['ms_per_step'],3))")
