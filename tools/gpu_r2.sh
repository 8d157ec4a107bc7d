#!/bin/bash
# One-B200 measurement pass of round 2 (run through gpurun): MUFU micro-benchmark, both bench arms, per-phase times of
# both algorithms, the ncu launch list of a node-grid step and one full capture of its kernels.  Outputs under
# gpurun_out/ with the prefix $TAG; the summaries that matter are copied to profiles/ by hand afterwards.
mkdir -p gpurun_out
T=${TAG:-r2}
( cd tools/ubench && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu 2>/dev/null && timeout 120 ./pipes ) > gpurun_out/${T}_ubench_pipes.txt 2>&1
timeout 900 python bench.py --steps ${STEPS:-100} --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?" >> gpurun_out/${T}_bench.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?" >> gpurun_out/${T}_bench_ref.err
timeout 200 python tools/phase_times.py --alg nodes --steps 40 > gpurun_out/${T}_phase_times.txt 2>&1
timeout 200 python tools/phase_times.py --alg direct --steps 40 >> gpurun_out/${T}_phase_times.txt 2>&1
timeout 200 python tools/prof_step.py --alg nodes --steps 2 --warmup 3 > gpurun_out/${T}_prof_plain.log 2>&1 && \
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${T}_launches_C4_nodes.csv python tools/prof_step.py --alg nodes --steps 2 --warmup 3 > gpurun_out/${T}_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --launch-skip 60 -c 16 -o gpurun_out/${T}_full_C4_nodes -f python tools/prof_step.py --alg nodes --steps 1 --warmup 3 > gpurun_out/${T}_ncu_full.log 2>&1
tail -n 2 gpurun_out/${T}_bench.err gpurun_out/${T}_bench_ref.err gpurun_out/${T}_ncu_full.log
cat gpurun_out/${T}_phase_times.txt
