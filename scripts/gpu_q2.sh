#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
echo cfg2; timeout 120 python scripts/prof_step.py --steps 3 --stages 2>&1 | grep '"step": 2' | grep -o '"ms": [0-9.]*\|"group": [0-9.]*' | paste - -
echo 5x; timeout 300 python scripts/prof_step.py --steps 2 --stages --nreads 5000000 --len 150 --M 3 --torchgen --seed 1003 --glen 150000000 2>&1 | grep '"step": 1\|checks' | grep -o '"ms": [0-9.]*\|"group": [0-9.]*\|"checks_ok": [a-z]*' | paste - - -
