#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
bash scripts/gpu_sweep.sh "--opt rebuild=2" "--opt rebuild=1" "--torchgen --nreads 5000000 --len 150 --M 3 --seed 1003 --opt rebuild=2" "--torchgen --nreads 5000000 --len 150 --M 3 --seed 1003 --opt rebuild=1"
