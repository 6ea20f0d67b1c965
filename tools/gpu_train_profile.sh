#!/bin/bash
# per-kernel ncu metrics of one eager G + D training step (batch 128): bash tools/gpu_train_profile.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_training_gpu.py tests/test_vgg_gpu.py -m gpu -q -x > gpurun_out/${tag}_train_pytest.log 2>&1; echo "pytest exit=$?"; tail -n 3 gpurun_out/${tag}_train_pytest.log
timeout 300 python tools/profile_train_step.py 128 > gpurun_out/${tag}_train_plain.log 2>&1 &&
timeout 1500 ncu --clock-control none --profile-from-start off \
  --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
  --csv --log-file gpurun_out/${tag}_train_kernel_metrics.csv python tools/profile_train_step.py 128 > gpurun_out/${tag}_train_ncu.log 2>&1
echo "metrics exit=$?"
python tools/train_kernel_table.py gpurun_out/${tag}_train_kernel_metrics.csv > gpurun_out/${tag}_train_kernel_table.txt; head -n 32 gpurun_out/${tag}_train_kernel_table.txt
