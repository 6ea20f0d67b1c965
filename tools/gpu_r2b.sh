#!/bin/bash
# per-kernel times of one 128-frame forward, standard vs high precision (ncu, base clocks not locked)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "wide or split" > gpurun_out/r02b_pytest.log 2>&1; echo "pytest exit=$?"; tail -n 3 gpurun_out/r02b_pytest.log
for prec in standard high; do
  timeout 300 python tools/profile_step.py --batch 128 --precision $prec > gpurun_out/r02b_plain_$prec.log 2>&1 &&
  timeout 900 ncu --clock-control none --profile-from-start off \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_bytes.sum,sm__cycles_elapsed.max,launch__grid_size,launch__registers_per_thread \
    --csv --log-file gpurun_out/r02b_kernel_metrics_$prec.csv python tools/profile_step.py --batch 128 --precision $prec > gpurun_out/r02b_ncu_$prec.log 2>&1
  echo "$prec metrics exit=$?"; cat gpurun_out/r02b_plain_$prec.log
  python tools/kernel_table.py gpurun_out/r02b_kernel_metrics_$prec.csv > gpurun_out/r02b_kernel_table_$prec.txt; tail -n 1 gpurun_out/r02b_kernel_table_$prec.txt
done
