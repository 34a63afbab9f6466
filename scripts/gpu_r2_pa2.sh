#!/bin/bash
# multiclass PA kernel: lane-prefetched features, 4 rows in flight; launch-bounds variants on config 4
O=gpurun_out/r2pa2; mkdir -p $O; rm -f $O/*.json
timeout -s KILL 900 python -m pytest tests/test_device_pa.py tests/test_device_dispatch.py -m gpu -x -q 2>&1 | tail -3
for mb in 4 3 2; do
FPS_PA_MINB=$mb timeout -s KILL 300 python benchmarks/run_configs.py --configs 4 --out $O/config4_minb$mb.json > $O/c4_$mb.log 2>&1
python -c "
import json
for l in open('$O/config4_minb$mb.json'):
    r=json.loads(l); print('minb', $mb, {k:r.get(k) for k in ('ms_per_step','without_limiter_ms_per_step','examples_per_sec','gather_scatter_gbs','error')})"
done
