#!/bin/bash
# round-2 first GPU session: new wide/split kernel tests, generator parity at 50 dB, precision A/B of the bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/r02a_smi.txt
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_generator_gpu.py -m gpu -q -s > gpurun_out/r02a_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r02a_pytest.log
grep -E "psnr|passed|failed|error|Error|rms error|u8:" gpurun_out/r02a_pytest.log | tail -n 40
for prec in standard high standard high; do
  timeout 600 python bench.py --no-cpu-baseline --no-train --precision $prec > gpurun_out/r02a_bench_$prec.json 2> gpurun_out/r02a_bench_$prec.err
  echo "bench $prec exit=$?"; tail -n 3 gpurun_out/r02a_bench_$prec.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r02a_bench_$prec.json"))
print("$prec", round(d["value"],1), "frames/s  e2e", round(d["e2e"]["value"],1), "conv TF/s", round(d["roofline"]["achieved"],1), d["clocks"], d["kernel_time_share"], d["gpu_launches"])
PY
done
