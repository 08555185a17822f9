#!/bin/bash
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/dp_check.py 2>&1 | grep -v "Setting OMP_NUM_THREADS\|^\*\*\*\*" | tail -40
