#!/bin/bash
# Small workloads (the reference's own shapes): single-chain bench lines and R independent chains on one GPU.
mkdir -p gpurun_out
T=${TAG:-r2k}
for w in C1 C2 C3; do
  python bench.py --workload $w --steps 400 --warmup 5 --no-pred --no-cpu > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench_$w.json").read().strip().splitlines()[-1])
print("$w", round(d["value"]), round(d["e2e"]["value"]), {k: round(v*1e3,1) for k,v in d["roofline"]["phase_ms_per_step"].items()}, d["gpu_launches"], d["clocks"])
PY
  for R in 8 32; do
    python bench.py --workload $w --steps 400 --warmup 5 --replicas $R > gpurun_out/${T}_bench_${w}_rep$R.json 2>> gpurun_out/${T}_bench_$w.err
    python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench_${w}_rep$R.json").read().strip().splitlines()[-1])
print("$w replicas", $R, round(d["value"]), round(d["single_chain_steps_per_sec"]), d["finite"])
PY
  done
done
tail -3 gpurun_out/${T}_bench_C2.err
