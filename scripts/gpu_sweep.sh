#!/bin/bash
# several staged profile steps with different options, one line of stages each
mkdir -p gpurun_out; : > gpurun_out/sweep.log
while [ $# -gt 0 ]; do
  echo "### $1" >> gpurun_out/sweep.log
  timeout 300 python scripts/prof_step.py --steps 2 --stages $1 2>&1 | grep -E "\"step\": 1|checks_ok|rror" >> gpurun_out/sweep.log
  shift
done
cat gpurun_out/sweep.log | cut -c1-400
