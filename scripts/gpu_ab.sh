#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
(cd ab_old && python build_native.py > /dev/null 2>&1)
nvidia-smi --query-gpu=name,clocks.max.mem,clocks.mem,power.limit,temperature.gpu --format=csv
for rep in 1 2; do
for d in ab_old .; do
  (cd $d && timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$d','ms',round(d['ms_per_step'],3),'Gr/s',round(d['value']/1e9,3))")
done; done
timeout 300 python -m pytest tests/test_device_mf.py -m gpu -q --timeout 150 -k combine 2>&1 | tail -12
