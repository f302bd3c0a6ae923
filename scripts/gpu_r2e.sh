#!/bin/bash
# round-2 call E: the unmodified driver test, the full-size CPU baseline record (three variants), the reference arm as the driver runs it
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_tools.py -m gpu -q > gpurun_out/r2e_tools.log 2>&1; echo "rc=$?" >> gpurun_out/r2e_tools.log; tail -4 gpurun_out/r2e_tools.log
BFQ_BENCH_EGAP_FULL=1 timeout 1500 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2e_ref_full.json 2> gpurun_out/r2e_ref_full.err; echo "ref rc=$?"
cat gpurun_out/r2e_ref_full.json | cut -c1-2500
