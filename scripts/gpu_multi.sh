#!/bin/bash
# gpu_multi.sh N : the bench as the driver launches it on N GPUs (NCCL_DEBUG=INFO, as the driver sets it) + the N-way PCIe probe
N=${1:-8}
mkdir -p gpurun_out
export NCCL_DEBUG=INFO
free -g | head -2 | tail -1; nproc; nvidia-smi topo -m 2>/dev/null | head -12 | cut -c1-140
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
grep -c "NCCL INFO" gpurun_out/bench_n$N.err; grep -m2 "Init COMPLETE" gpurun_out/bench_n$N.err | cut -c1-160; tail -3 gpurun_out/bench_n$N.err | cut -c1-300
unset NCCL_DEBUG
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 scripts/pcie_probe_multi.py > gpurun_out/pcie_n$N.json 2> gpurun_out/pcie_n$N.err; echo "probe rc=$?"; cat gpurun_out/pcie_n$N.json
python - <<PY
import json
d=json.load(open("gpurun_out/bench_n$N.json"))
print("value", round(d["value"],2), d["scaling"], "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"],2), round(d["e2e"]["ms_per_step"],1), "checks", d.get("checks_ok"), d["e2e"]["equals_device_result"])
w=d.get("weak_config2"); print("weak cfg2", round(w["value"],2), round(w["ms_per_step"],2), "e2e", round(w["e2e"]["value"],2)) if w else None
s=d.get("single_collection"); print("single", round(s["value"],2), round(s["ms_per_step"],1), s["checks_ok"], s["wide_positions"], s["stage_ms_per_step"], s["kernel_class_ms_per_step"]) if s else None
PY
