mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
GPTB_PIPE_GROUPS=0 timeout 300 python tools/phase_times.py --steps 100 2>&1 | grep "unprofiled" | sed "s/^/G0 /"
timeout 300 python tools/phase_times.py --steps 100 2>&1 | grep "unprofiled" | sed "s/^/auto /"
N=2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 200 --warmup 5 > gpurun_out/r1k_bench_n$N.json 2> gpurun_out/r1k_bench_n$N.err; echo "rc=$?"
wc -l gpurun_out/r1k_bench_n$N.json; head -c 200 gpurun_out/r1k_bench_n$N.json
