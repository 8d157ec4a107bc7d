mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
for w in C1 C2 C3; do
  timeout 600 python bench.py --workload $w --no-pred --no-cpu --steps 300 --warmup 10 > gpurun_out/r1j_bench_$w.json 2> gpurun_out/r1j_bench_$w.err; echo "$w rc=$?"
done
