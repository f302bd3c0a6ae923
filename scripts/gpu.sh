#!/bin/bash
# The GPU-box tasks of this repository, one script (run through gpurun: `gpurun -- 'bash scripts/gpu.sh <task> [args]'`).
# Everything a task writes goes to gpurun_out/ (scratch); summaries meant to be kept are copied to profiles/ by hand.
#   tests                 pytest -m gpu (whole suite)
#   stage [prof_step args] parity of part A / pipeline / dist, then per-stage CUDA-event times at config 2 and on a config-4 block
#   ab OPTION V1 V2 ...   scripts/ab_option.py on config 2 and on a config-4 block (outputs must be bit-identical)
#   bench [N]             bench.py as the driver launches it on N GPUs (default 1) + the reference arm at N = 1
#   multi N               bench.py on N GPUs + the N-way concurrent PCIe probe
#   launches              ncu launch list (gpu__time_duration.sum, cold-cache, serialised) of one bench step -> launches.csv
#   ncu REGEX NAME [prof_step args]   one `ncu --set full --import-source on` capture of the kernels matching REGEX
#   cpu-baseline          the reference arm with the full-size eGap leg (minutes of CPU; recorded once per pool)
set -u
mkdir -p gpurun_out
task=${1:-tests}; shift || true
CFG4="--nreads 5000000 --len 150 --M 3 --torchgen --glen 150000000 --seed 1003"
case "$task" in
  tests)
    timeout 1400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log; tail -5 gpurun_out/pytest_gpu.log ;;
  stage)
    timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_dist.py -m gpu -x -q > gpurun_out/stage_parity.log 2>&1; echo "rc=$?" >> gpurun_out/stage_parity.log; tail -2 gpurun_out/stage_parity.log
    echo "== cfg2"; timeout 300 python scripts/prof_step.py --steps 4 --stages "$@" 2>&1 | grep '"step": 3'
    echo "== cfg4 block"; timeout 300 python scripts/prof_step.py $CFG4 --steps 3 --stages "$@" 2>&1 | grep '"step": 2\|checks_ok' | cut -c1-600 ;;
  ab)
    timeout 300 python scripts/ab_option.py "$@"; timeout 400 python scripts/ab_option.py "$@" --cfg4 ;;
  bench)
    N=${1:-1}
    if [ "$N" = 1 ]; then
      timeout 1200 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
      timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref_n1.err; echo "reference rc=$?"
      cut -c1-700 gpurun_out/bench_n1.json; cut -c1-500 gpurun_out/bench_ref_n1.json
    else
      bash scripts/gpu_multi.sh "$N"
    fi ;;
  multi)
    bash scripts/gpu_multi.sh "${1:-8}" ;;
  launches)
    timeout 300 python scripts/prof_step.py --steps 1 "$@" > gpurun_out/ncu_plain.log 2>&1 || { tail -5 gpurun_out/ncu_plain.log; exit 1; }
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python scripts/prof_step.py --steps 1 "$@" > gpurun_out/ncu_launches.log 2>&1
    echo "rc=$?"; wc -l gpurun_out/launches.csv ;;
  ncu)
    bash scripts/gpu_ncu_one.sh "$@" ;;
  cpu-baseline)
    BFQ_BENCH_EGAP_FULL=1 timeout 1500 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/ref_full.json 2> gpurun_out/ref_full.err; echo "rc=$?"; cut -c1-2500 gpurun_out/ref_full.json ;;
  *)
    echo "unknown task $task"; exit 2 ;;
esac
