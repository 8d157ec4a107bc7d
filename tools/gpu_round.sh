mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 15
timeout 600 python tools/nt_time.py > gpurun_out/r1i_fulltheta_C3.jsonl 2> gpurun_out/r1i_fulltheta.err; echo "nt rc=$?"; cat gpurun_out/r1i_fulltheta_C3.jsonl; tail -n 5 gpurun_out/r1i_fulltheta.err
