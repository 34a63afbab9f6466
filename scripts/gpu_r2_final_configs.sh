#!/bin/bash
# final one-GPU snapshot of every config of benchmarks/run_configs.py
O=gpurun_out/r2finalcfg; mkdir -p $O; rm -f $O/configs.jsonl
timeout 1200 python benchmarks/run_configs.py --configs 1,2,3,4,6,7,8 --out $O/configs.jsonl > $O/configs.log 2>&1; echo "rc=$?"
python -c "
import json
for l in open('$O/configs.jsonl'):
    r=json.loads(l); c=r.get('config')
    keys={1:['device_ratings_per_sec','pull_limit'],2:['ratings_per_sec','nan_count','pull_limit'],3:['examples_per_sec','without_limiter'],4:['examples_per_sec','without_limiter_ms_per_step','ms_per_step'],6:['packed_pipeline_all_tiles_ms','packed_pipeline_pruned_ms'],7:['device_batch8192','device_batch256'],8:['records_per_sec']}.get(c,[])
    print(c, {k:(r.get(k) if not isinstance(r.get(k),dict) else {kk:vv for kk,vv in r[k].items() if kk in ('events_per_sec','ms_per_step','examples_per_sec')}) for k in keys}, r.get('error'))"
