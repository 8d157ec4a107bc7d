#!/bin/bash
# Final measurements of the round on one B200: bench line, ncu launch list + full capture of the reduced posterior sweep.
mkdir -p gpurun_out
T=${TAG:-r2n}
timeout 600 python bench.py --steps 100 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 120 python tools/prof_pred.py > gpurun_out/${T}_prof_pred_plain.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${T}_launches_C5_pred.csv python tools/prof_pred.py > gpurun_out/${T}_ncu_pred_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"interp_kernel|contract_fwd3_kernel" -c 4 -o gpurun_out/${T}_full_pred -f python tools/prof_pred.py > gpurun_out/${T}_ncu_pred_full.log 2>&1
tail -2 gpurun_out/${T}_ncu_pred_full.log
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["e2e"]["blocking_api_value"], d["e2e"]["stored_chain"]["value"], d["pred"]["value"], d["roofline"]["frac"], d["roofline"]["step"]["frac"], d["clocks"])
PY
