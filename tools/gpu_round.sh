mkdir -p gpurun_out
GPTB_PIPE_GROUPS=0 timeout 300 python tools/phase_times.py --steps 60 2>&1 | grep "ms/step" | sed "s/^/base /"
for v in acc128; do
  GPTB_PIPE_GROUPS=0 GPTB200_LIB=$PWD/gpt_b200/lib/v_$v/libgptb200.so timeout 300 python tools/phase_times.py --steps 60 2>&1 | grep "ms/step" | sed "s/^/$v /"
done
