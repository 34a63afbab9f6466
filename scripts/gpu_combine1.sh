#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
timeout 300 python -m pytest tests/test_device_mf.py -m gpu -q --timeout 150 2>&1 | tail -12
for v in direct combine; do for d in movielens uniform; do
  timeout 300 python bench.py --steps 20 --warmup 3 --variant $v --item-dist $d > gpurun_out/bench_${v}_${d}_1.log 2>&1
  tail -1 gpurun_out/bench_${v}_${d}_1.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v','$d','ms',round(d['ms_per_step'],3),'Gr/s',round(d['value']/1e9,3),'e2e',round(d['e2e']['value']/1e9,3))"
done; done
