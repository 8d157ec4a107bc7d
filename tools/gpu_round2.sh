#!/bin/bash
# Multi-GPU bench line on NG GPUs of one box (run through `gpurun --gpus NG`), launched as the driver launches it.
mkdir -p gpurun_out
N=${NG:-2}
T=${TAG:-r1}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err; echo "rc=$?" >> gpurun_out/${T}_bench_n$N.err
tail -n 3 gpurun_out/${T}_bench_n$N.err; wc -l gpurun_out/${T}_bench_n$N.json; head -c 300 gpurun_out/${T}_bench_n$N.json
