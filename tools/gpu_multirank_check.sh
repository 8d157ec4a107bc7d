#!/bin/bash
# on 2 GPUs of one box: parity tests on GPU 0, then the row-sharded chain against the one-rank chain
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 tools/check_multirank_host.py 2>&1 | grep -v "OMP_NUM\|\*\*\*\*\|^$" | tail -n 6
