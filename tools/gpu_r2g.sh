#!/bin/bash
# TMA x tile in the fused MAT epilogue + fused state encoder: kernel tests, generator parity, bench A/B, kernel table
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_generator_gpu.py -m gpu -q -s > gpurun_out/r02g_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/r02g_pytest.log
grep -E "psnr|passed|failed|error|Error|rms error|u8:|rel rms" gpurun_out/r02g_pytest.log | tail -n 40
for prec in standard high standard high; do
  timeout 600 python bench.py --no-cpu-baseline --no-train --precision $prec > gpurun_out/r02g_bench_$prec.json 2> gpurun_out/r02g_bench_$prec.err
  echo "bench $prec exit=$?"; tail -n 3 gpurun_out/r02g_bench_$prec.err
  python - <<PY
import json
d=json.load(open("gpurun_out/r02g_bench_$prec.json"))
print("$prec", round(d["value"],1), "frames/s  e2e", round(d["e2e"]["value"],1), "conv TF/s", round(d["roofline"]["achieved"],1), d["clocks"], d["kernel_time_share"], d["gpu_launches"])
PY
done
for prec in standard high; do
  timeout 300 python tools/profile_step.py --batch 128 --precision $prec > gpurun_out/r02g_plain_$prec.log 2>&1 &&
  timeout 900 ncu --clock-control none --profile-from-start off \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_bytes.sum,sm__cycles_elapsed.max,launch__grid_size,launch__registers_per_thread \
    --csv --log-file gpurun_out/r02g_kernel_metrics_$prec.csv python tools/profile_step.py --batch 128 --precision $prec > gpurun_out/r02g_ncu_$prec.log 2>&1
  echo "$prec metrics exit=$?"; cat gpurun_out/r02g_plain_$prec.log
  python tools/kernel_table.py gpurun_out/r02g_kernel_metrics_$prec.csv > gpurun_out/r02g_kernel_table_$prec.txt; tail -n 1 gpurun_out/r02g_kernel_table_$prec.txt
done
