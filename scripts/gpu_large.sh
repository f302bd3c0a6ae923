#!/bin/bash
# device-resident steps of the larger BASELINE configurations (reads generated on the device, size-independent checks)
mkdir -p gpurun_out
echo "cfg3: 10M x 150 bp, 40x, M=1 B=1"
timeout 600 python scripts/prof_step.py --steps 2 --stages --nreads 10000000 --len 150 --M 1 --B 1 --torchgen --seed 1002 2>&1 | tail -3 | cut -c1-700
echo "cfg4 block: 5M x 150 bp of a 150 Mbp genome (5x per block), M=3"
timeout 600 python scripts/prof_step.py --steps 2 --stages --nreads 5000000 --len 150 --M 3 --torchgen --seed 1003 --glen 150000000 2>&1 | tail -3 | cut -c1-700
echo "5M x 150 bp at 40x, M=3"
timeout 600 python scripts/prof_step.py --steps 2 --stages --nreads 5000000 --len 150 --M 3 --torchgen --seed 1003 2>&1 | tail -3 | cut -c1-700
