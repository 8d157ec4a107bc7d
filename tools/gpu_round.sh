mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r1h_bench2.json 2> gpurun_out/r1h_bench2.err; echo "bench rc=$?"
nproc; uptime
