#!/bin/bash
# source-level ncu capture of one kernel class: gpu_ncu_one.sh <kernel regex> <out name> [prof_step args]
# (the plain command runs first: a number printed under ncu is never a bench value)
mkdir -p gpurun_out
K="$1"; O="$2"; shift 2
timeout 300 python scripts/prof_step.py --steps 1 "$@" > gpurun_out/ncu_plain.log 2>&1 || { tail -5 gpurun_out/ncu_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$K" -c ${NCU_COUNT:-1} -f -o gpurun_out/$O python scripts/prof_step.py --steps 1 "$@" > gpurun_out/ncu_$O.log 2>&1
tail -2 gpurun_out/ncu_$O.log | cut -c1-200; ls -la gpurun_out/$O.ncu-rep
