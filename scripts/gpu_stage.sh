#!/bin/bash
# parity of part A / pipeline, then stage times at config 2 and on a config-4 block (5M x 150 bp of the 150 Mbp genome, M=3)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_dist.py -m gpu -x -q > gpurun_out/stage_parity.log 2>&1; echo rc=$? >> gpurun_out/stage_parity.log; tail -2 gpurun_out/stage_parity.log
echo "== cfg2"; timeout 300 python scripts/prof_step.py --steps 4 --stages "$@" 2>&1 | grep '"step": 3'
echo "== cfg4 block"; timeout 300 python scripts/prof_step.py --nreads 5000000 --len 150 --M 3 --torchgen --glen 150000000 --seed 1003 --steps 3 --stages "$@" 2>&1 | grep '"step": 2\|checks_ok' | cut -c1-600
