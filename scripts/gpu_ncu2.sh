#!/bin/bash
# usage: gpu_ncu2.sh "<regex1> <regex2> ..." : one full capture (first launch) per kernel regex
mkdir -p gpurun_out
python scripts/prof_step.py --steps 1 > gpurun_out/ncu_plain.log 2>&1 || exit 1
for K in $1; do
  ncu --set full --clock-control none --import-source on -k regex:$K -c 1 -o gpurun_out/prof_$K -f \
      python scripts/prof_step.py --steps 1 > gpurun_out/ncu_$K.log 2>&1
  tail -1 gpurun_out/ncu_$K.log
done
