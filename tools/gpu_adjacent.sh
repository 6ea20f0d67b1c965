#!/bin/bash
# ncu evidence for the adjacent components: launch list + full capture.  usage (under gpurun): bash tools/gpu_adjacent.sh <tag>
tag=${1:-r01adj}
mkdir -p gpurun_out
timeout 300 python tools/profile_adjacent.py > gpurun_out/${tag}_plain.log 2>&1 || { cat gpurun_out/${tag}_plain.log; exit 1; }
cat gpurun_out/${tag}_plain.log
timeout 600 ncu --clock-control none --profile-from-start off \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_fma.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_elapsed.max,launch__grid_size,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active \
    --csv --log-file gpurun_out/${tag}_kernel_metrics.csv python tools/profile_adjacent.py > gpurun_out/${tag}_ncu_metrics.log 2>&1
echo "metrics exit=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:'ensemble_transition|window_gather|slac_conv1' -o gpurun_out/${tag}_prof -f python tools/profile_adjacent.py > gpurun_out/${tag}_ncu_full.log 2>&1
echo "full capture exit=$?"; tail -n 2 gpurun_out/${tag}_ncu_full.log
ls -la gpurun_out | grep ${tag}
