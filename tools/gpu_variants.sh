#!/bin/bash
# Times the step phases with each experiment build of the library under build_dbg/<name>/ (see build.sh).
mkdir -p gpurun_out
T=${TAG:-var}
for d in build_dbg/*/; do
  n=$(basename $d)
  [ -f $d/libgptb200.so ] || continue
  echo "== $n" >> gpurun_out/${T}_variants.txt
  GPTB200_LIB=$PWD/$d/libgptb200.so timeout 200 python tools/phase_times.py --alg ${ALG:-nodes} --steps 40 2>&1 | grep "ms/step" >> gpurun_out/${T}_variants.txt
done
cat gpurun_out/${T}_variants.txt
