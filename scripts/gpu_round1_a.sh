#!/bin/bash
# first GPU call of the session: parity tests, bench, ncu launch list + full capture of the radix pass
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python scripts/prof_step.py --steps 2 --stages > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python scripts/prof_step.py --steps 2 > gpurun_out/ncu_launches.log 2>&1
python scripts/prof_step.py --steps 2 > gpurun_out/prof_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rs_onesweep -s 8 -c 3 -o gpurun_out/prof_onesweep \
    python scripts/prof_step.py --steps 2 > gpurun_out/ncu_full.log 2>&1
python scripts/prof_step.py --steps 2 > gpurun_out/prof_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:regroup_kernel -s 1 -c 2 -o gpurun_out/prof_regroup \
    python scripts/prof_step.py --steps 2 > gpurun_out/ncu_full2.log 2>&1
ls -la gpurun_out
