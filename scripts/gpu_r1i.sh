#!/bin/bash
# round-1i: parity tests, staged step, source-level ncu capture of group_tile_kernel and one onesweep pass
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python scripts/prof_step.py --steps 3 --stages > gpurun_out/prof_plain.log 2>&1; echo "prof rc=$?"
tail -25 gpurun_out/prof_plain.log
timeout 600 ncu --set full --clock-control none --import-source on \
    -k regex:'group_tile_kernel|rs_onesweep_kernel' -c 3 -f -o gpurun_out/prof_r1i python scripts/prof_step.py --steps 1 > gpurun_out/ncu_r1i.log 2>&1
tail -2 gpurun_out/ncu_r1i.log | cut -c1-200; ls -la gpurun_out/prof_r1i.ncu-rep
