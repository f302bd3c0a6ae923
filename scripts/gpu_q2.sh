#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
echo cfg2-flat; timeout 120 python scripts/prof_step.py --steps 3 --stages --opt rebuild=1 2>&1 | grep '"step": 2' | cut -c1-500
echo 5x; timeout 300 python scripts/prof_step.py --steps 2 --stages --nreads 5000000 --len 150 --M 3 --torchgen --seed 1003 --glen 150000000 2>&1 | grep '"step": 1\|checks' | cut -c1-500
echo cfg3; timeout 300 python scripts/prof_step.py --steps 2 --stages --nreads 10000000 --len 150 --M 1 --B 1 --torchgen --seed 1002 2>&1 | grep '"step": 1\|checks' | cut -c1-500
