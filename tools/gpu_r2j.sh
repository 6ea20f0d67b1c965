#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --workload train --batch 256 --steps 5 --warmup 3 > gpurun_out/r02j_train256.json 2> gpurun_out/r02j_train256.err; echo "train256 exit=$?"; cut -c1-900 gpurun_out/r02j_train256.json
timeout 600 python bench.py --workload train --batch 16 --steps 10 --warmup 3 > gpurun_out/r02j_train16.json 2> gpurun_out/r02j_train16.err; echo "train16 exit=$?"; cut -c1-600 gpurun_out/r02j_train16.json
