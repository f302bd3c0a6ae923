#!/bin/bash
mkdir -p gpurun_out
for d in 0 1 2 3 4 8 12; do
echo "dbg=$d"; BFQ_DBG_GROUP=$d timeout 120 python scripts/prof_step.py --steps 3 --stages 2>&1 | grep '"step": 2' | grep -o '"group": [0-9.]*'
done
