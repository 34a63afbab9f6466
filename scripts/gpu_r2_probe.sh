#!/bin/bash
N=${1:-2}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 scripts/p2p_symmetric_probe.py 2>&1 | grep -v Warning | tail -8
