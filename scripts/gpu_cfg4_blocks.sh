#!/bin/bash
# BASELINE config 4 in block mode: 8 blocks of 5M x 150 bp reads of the 150 Mbp genome (5x per block), M=3, one block per GPU,
# all GPUs at once (no collective on the data path).  Reads are generated on each device; size-independent checks per block.
N=${1:-8}
mkdir -p gpurun_out
for i in $(seq 0 $((N - 1))); do
  CUDA_VISIBLE_DEVICES=$i timeout 500 python scripts/prof_step.py --steps 3 --nreads 5000000 --len 150 --M 3 --torchgen --seed $((1003 + i)) --glen 150000000 \
      > gpurun_out/cfg4_block$i.log 2>&1 &
done
wait
for i in $(seq 0 $((N - 1))); do echo "block $i:"; grep '"step"\|checks_ok' gpurun_out/cfg4_block$i.log | cut -c1-160; done
