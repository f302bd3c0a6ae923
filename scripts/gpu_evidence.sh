#!/bin/bash
# round-1 evidence (latest: r1j): parity tests, both bench arms, ncu launch list, one --set full capture of the hot kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -2 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -2 gpurun_out/bench.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python scripts/prof_step.py --steps 2 > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python scripts/prof_step.py --steps 2 > gpurun_out/ncu_launches.log 2>&1
python scripts/prof_step.py --steps 1 > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:'group_tile_kernel|rs_onesweep_kernel|walk_kernel|cluster_tile_kernel|rs_suffix_keys_kernel|pack_reads_kernel|pack3_kernel|rank_' \
    -c 16 -f -o gpurun_out/prof_r1j_all python scripts/prof_step.py --steps 1 > gpurun_out/ncu_all.log 2>&1
tail -1 gpurun_out/ncu_all.log | cut -c1-200; ls -la gpurun_out/*.ncu-rep
