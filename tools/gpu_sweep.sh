#!/bin/bash
# sweep of the dimension-group pipeline's knobs (GPTB_PIPE_GROUPS / GPTB_PIPE_RESERVE) at C4
for cfg in "0 0" "2 3" "2 4" "2 5" "2 6"; do
  set -- $cfg
  GPTB_PIPE_GROUPS=$1 GPTB_PIPE_RESERVE=$2 timeout 300 python tools/phase_times.py --steps 100 2>&1 | grep unprofiled | sed "s/^/G=$1 R=$2 /"
done
