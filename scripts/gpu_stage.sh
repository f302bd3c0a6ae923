#!/bin/bash
# quick look: staged per-class times of a device-resident step (extra args go to prof_step.py)
mkdir -p gpurun_out
timeout 300 python scripts/prof_step.py --steps 3 --stages "$@" > gpurun_out/prof_plain.log 2>&1; echo "prof rc=$?"
tail -3 gpurun_out/prof_plain.log | cut -c1-600
