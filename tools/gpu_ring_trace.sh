#!/bin/bash
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 tools/ring_trace.py 2>&1 | grep "^rank" | sort
