#!/bin/bash
# Round-end style check on ONE B200 (run through gpurun): parity tests, smoke, both bench arms, the ncu launch list and
# one full capture of a step.  TAG names the outputs under gpurun_out/.
mkdir -p gpurun_out
T=${TAG:-r1}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${T}_smoke.log
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?" >> gpurun_out/${T}_bench.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?" >> gpurun_out/${T}_bench_ref.err
timeout 300 python tools/prof_step.py --steps 2 --warmup 3 > gpurun_out/${T}_prof_plain.log 2>&1 && \
GPTB_PIPE_GROUPS=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_C4.csv python tools/prof_step.py --steps 2 --warmup 3 > gpurun_out/${T}_ncu_launch.log 2>&1
GPTB_PIPE_GROUPS=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fwd_tc|bwd_tc|contract_tc|k_noise|k_prepare|k_small|k_apply|gather_kernel' --launch-skip 24 -c 8 -o gpurun_out/${T}_full_C4 -f python tools/prof_step.py --steps 1 --warmup 3 > gpurun_out/${T}_ncu_full.log 2>&1
tail -n 3 gpurun_out/${T}_pytest.log gpurun_out/${T}_smoke.log gpurun_out/${T}_bench.err gpurun_out/${T}_bench_ref.err gpurun_out/${T}_ncu_full.log
