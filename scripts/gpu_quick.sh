#!/bin/bash
# quick GPU check: parity tests then a staged profile step; extra args go to prof_step.py
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python scripts/prof_step.py --steps 3 --stages "$@" > gpurun_out/prof_plain.log 2>&1; echo "prof rc=$?"
cat gpurun_out/prof_plain.log | tail -5
