#!/bin/bash
# runs the bring-up probe groups, each in its own process under a timeout
mkdir -p gpurun_out
for g in "$@"; do
  timeout 240 python tools/gpu_probe.py $g > gpurun_out/probe_$g.log 2>&1
  echo "group $g exit=$?" | tee -a gpurun_out/probe_$g.log
  tail -n 60 gpurun_out/probe_$g.log
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
