mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/g4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g4_pytest.log
timeout 300 python tools/phase_times.py --steps 50 > gpurun_out/g4_phase_base.log 2>&1
for v in w20 w28; do
  GPTB200_LIB=$PWD/gpt_b200/lib/v_$v/libgptb200.so timeout 300 python tools/phase_times.py --steps 50 > gpurun_out/g4_phase_$v.log 2>&1
done
tail -n 3 gpurun_out/g4_*.log
