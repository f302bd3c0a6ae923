#!/bin/bash
# usage: gpu_ncu.sh <kernel regex> <skip> <count> <outname> [prof_step args...]
K=$1; S=$2; C=$3; O=$4; shift 4
mkdir -p gpurun_out
python scripts/prof_step.py --steps 1 "$@" > gpurun_out/ncu_plain_$O.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c $C -o gpurun_out/prof_$O \
    python scripts/prof_step.py --steps 1 "$@" > gpurun_out/ncu_$O.log 2>&1
tail -3 gpurun_out/ncu_$O.log
