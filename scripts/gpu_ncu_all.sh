#!/bin/bash
# One ncu --set full invocation over the hot kernels of one device-resident step (first launch of each class is enough:
# 6 onesweep passes + group + cluster + walk + keys + pack).  Runs the plain command first.
mkdir -p gpurun_out
python scripts/prof_step.py --steps 1 > gpurun_out/ncu_plain.log 2>&1 || { tail -5 gpurun_out/ncu_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on \
    -k regex:'group_tile_kernel|rs_onesweep_kernel|walk_kernel|cluster_tile_kernel|rs_suffix_keys_kernel|pack_reads_kernel|pack3_kernel' \
    -c 16 -f -o gpurun_out/prof_r1g_all python scripts/prof_step.py --steps 1 > gpurun_out/ncu_all.log 2>&1
tail -2 gpurun_out/ncu_all.log | cut -c1-200; ls -la gpurun_out/prof_r1g_all.ncu-rep
