#!/bin/bash
mkdir -p gpurun_out
python build_native.py > gpurun_out/build.log 2>&1 || { tail -30 gpurun_out/build.log; exit 1; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $TR scripts/p2p_random_probe.py > gpurun_out/p2p_random.log 2>&1; echo "probe exit $?"; grep -E "GB/s|rror" gpurun_out/p2p_random.log | tail -70
timeout 200 python -m pytest tests/test_device_mf.py -m gpu -q --timeout 100 2>&1 | tail -3
for u in 1 2 4; do for b in 4 6 8; do
  timeout 200 python bench.py --steps 10 --warmup 3 --no-e2e --users 20000000 --items 10000000 --unroll $u --blocks-per-sm $b 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('unroll',$u,'bps',$b,'ms',round(d['ms_per_step'],3),'Gr/s',round(d['value']/1e9,3))"
done; done
