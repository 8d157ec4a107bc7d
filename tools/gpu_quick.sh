#!/bin/bash
# quick check on one B200: parity tests, per-phase times of the one-stream step, unprofiled step time, bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 6
GPTB_PIPE_GROUPS=0 timeout 300 python tools/phase_times.py --steps 100 2>&1 | grep "unprofiled" | sed "s/^/G0 /"
timeout 300 python tools/phase_times.py --steps 100 2>&1 | grep "unprofiled" | sed "s/^/auto /"
timeout 600 python bench.py --no-cpu > gpurun_out/quick_bench.json 2> gpurun_out/quick_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/quick_bench.json')); print('value', d['value'], 'e2e', d['e2e']['value'], d['e2e'].get('blocking_api_value'), 'pred', d['pred']['value'])"
