#!/bin/bash
O=gpurun_out/r2graph; mkdir -p $O; rm -f $O/*.jsonl
timeout -s KILL 900 python benchmarks/run_configs.py --configs 7 --out $O/config7.jsonl > $O/c7.log 2>&1; echo "rc=$?"; python -c "
import json
r=json.loads(open('$O/config7.jsonl').read().splitlines()[-1])
for k,v in r.items():
    if isinstance(v,dict) and 'events_per_sec' in v: print(k, round(v['events_per_sec']), 'ev/s ndcg', v.get('ndcg_last20k', v.get('ndcg')), 'hit', v.get('hit_rate_last20k', v.get('hit_rate')))
print(r.get('error'))"
