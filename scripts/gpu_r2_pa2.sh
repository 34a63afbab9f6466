#!/bin/bash
O=gpurun_out/r2pa2; mkdir -p $O; rm -f $O/*.json
for th in 256 128 64; do
FPS_PA_THREADS=$th timeout -s KILL 300 python benchmarks/run_configs.py --configs 4 --out $O/config4_th$th.json > $O/c4_$th.log 2>&1
python -c "
import json
for l in open('$O/config4_th$th.json'):
    r=json.loads(l); print('threads', $th, {k:r.get(k) for k in ('ms_per_step','without_limiter_ms_per_step','examples_per_sec','error')})"
done
FPS_PA_THREADS=128 timeout -s KILL 600 python -m pytest tests/test_device_pa.py tests/test_device_dispatch.py -m gpu -x -q 2>&1 | tail -2
