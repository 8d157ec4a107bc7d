mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 200 --warmup 5 > gpurun_out/g7_bench_n2.json 2> gpurun_out/g7_bench_n2.err; echo "rc=$?" >> gpurun_out/g7_bench_n2.err
tail -n 5 gpurun_out/g7_bench_n2.err; head -c 1500 gpurun_out/g7_bench_n2.json
