#!/bin/bash
# full-theta comparator: parity tests and the config-3 timing (tools/nt_time.py)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "fulltheta" 2>&1 | tail -n 3
timeout 600 python tools/nt_time.py > gpurun_out/${TAG:-r1m}_fulltheta_C3.jsonl 2> gpurun_out/${TAG:-r1m}_fulltheta.err; echo "nt rc=$?"
cat gpurun_out/${TAG:-r1m}_fulltheta_C3.jsonl | cut -c1-700; tail -n 3 gpurun_out/${TAG:-r1m}_fulltheta.err
