#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for b in 64; do echo "cfg2 batch_min_groups=$b"; timeout 120 python scripts/prof_step.py --steps 3 --stages --opt batch_min_groups=$b 2>&1 | grep '"step": 2' | grep -o '"group": [0-9.]*'; done
for b in 64 1000000; do echo "5x batch_min_groups=$b"; timeout 300 python scripts/prof_step.py --steps 2 --stages --nreads 5000000 --len 150 --M 3 --torchgen --seed 1003 --glen 150000000 --opt batch_min_groups=$b 2>&1 | grep '"step": 1\|checks' | grep -o '"group": [0-9.]*\|"checks_ok": [a-z]*'; done
