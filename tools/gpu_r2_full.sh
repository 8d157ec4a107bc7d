#!/bin/bash
# Round-2 validation pass on one B200: the whole -m gpu suite, smoke(), then tools/gpu_r2.sh (bench arms, phases, ncu).
mkdir -p gpurun_out
T=${TAG:-r2}
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${T}_smoke.log
timeout 300 python tools/pred_time.py 200 1000000 nodes > gpurun_out/${T}_pred_time.txt 2>&1
tail -n 3 gpurun_out/${T}_pytest_gpu.log gpurun_out/${T}_smoke.log gpurun_out/${T}_pred_time.txt
TAG=$T bash tools/gpu_r2.sh
