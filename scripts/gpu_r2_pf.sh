#!/bin/bash
O=gpurun_out/r2pf; mkdir -p $O
timeout -s KILL 600 python -m pytest tests/test_device_mf.py -m gpu -x -q 2>&1 | tail -2
for lim in 1600 16384; do for pf in 0 1; do
echo "limit $lim prefetch $pf: $(FPS_MF_PREFETCH=$pf timeout -s KILL 300 python bench.py --steps 20 --warmup 5 --pull-limit $lim --baseline-steps 0 --no-e2e 2>/dev/null | python -c 'import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d["ms_per_step"],3), "ms", round(d["value"]/1e9,3), "G/s maxif", d.get("max_in_flight"), "nan", d.get("nan_count"))')"
done; done
timeout -s KILL 300 python benchmarks/run_configs.py --configs 1 --out $O/config1.jsonl > $O/c1.log 2>&1; tail -1 $O/config1.jsonl | cut -c1-300
