#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02h_bench.json 2> gpurun_out/r02h_bench.err; echo "bench exit=$?"; tail -n 5 gpurun_out/r02h_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/r02h_bench.json"))
for k in ("value","ms_per_step","e2e","gpu_launches","clocks","roofline","roofline_norm","kernel_time_share","parity_check","latency_b16","eager_gpu_baseline","training_step","rollout","cpu_baseline"):
    print(k, json.dumps(d.get(k))[:700])
PY
S2P_CONV_CLUSTER=1 python bench.py --steps 1 --warmup 1 --no-train --no-cpu-baseline 2>&1 | tail -n 2
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 | cut -c1-400
