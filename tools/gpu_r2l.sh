#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_generator_gpu.py -m gpu -q -x > gpurun_out/r02l_pytest.log 2>&1; echo "pytest exit=$?"; tail -n 12 gpurun_out/r02l_pytest.log
for prec in high; do
  timeout 300 python tools/profile_step.py --batch 128 --precision $prec > gpurun_out/r02l_plain_$prec.log 2>&1 &&
  timeout 900 ncu --clock-control none --profile-from-start off \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_bytes.sum,sm__cycles_elapsed.max,launch__grid_size,launch__registers_per_thread \
    --csv --log-file gpurun_out/r02l_kernel_metrics_$prec.csv python tools/profile_step.py --batch 128 --precision $prec > gpurun_out/r02l_ncu_$prec.log 2>&1
  echo "$prec metrics exit=$?"; cat gpurun_out/r02l_plain_$prec.log
  python tools/kernel_table.py gpurun_out/r02l_kernel_metrics_$prec.csv > gpurun_out/r02l_kernel_table_$prec.txt; tail -n 12 gpurun_out/r02l_kernel_table_$prec.txt | cut -c1-100
done
for i in 1 2; do timeout 600 python bench.py --no-cpu-baseline --no-train 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), d['roofline']['frac'], d['clocks']['sm_mhz'], d['parity_check']['psnr_db'])"; done
