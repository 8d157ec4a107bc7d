timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
GPTB_PIPE_GROUPS=0 timeout 300 python tools/phase_times.py --steps 100 2>&1 | grep "ms/step" | sed "s/^/G0 /"
timeout 300 python tools/phase_times.py --steps 100 2>&1 | grep "unprofiled" | sed "s/^/auto /"
