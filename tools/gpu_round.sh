#!/bin/bash
# One GPU-box session: parity tests, bench, ncu launch list and a bounded full capture.
# usage (under gpurun): bash tools/gpu_round.sh <tag> [full_capture_skip] [full_capture_count]
tag=${1:-r01}
skip=${2:-20}
cnt=${3:-4}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/${tag}_smi.txt
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit=$?" | tee -a gpurun_out/${tag}_pytest.log
tail -n 25 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit=$?"
cat gpurun_out/${tag}_bench.json; tail -n 5 gpurun_out/${tag}_bench.err
timeout 300 python tools/profile_step.py --batch 128 > gpurun_out/${tag}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/${tag}_launches.csv python tools/profile_step.py --batch 128 > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "launch list exit=$?"; cat gpurun_out/${tag}_plain.log
# compact metrics for every conv / norm launch of the forward
timeout 900 ncu --clock-control none --profile-from-start off -k regex:'conv_igemm|mat_norm|stem|torgb' \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tensor.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__m_xbar2l1tex_read_bytes.sum,sm__cycles_elapsed.max,launch__grid_size,launch__registers_per_thread \
    --csv --log-file gpurun_out/${tag}_kernel_metrics.csv python tools/profile_step.py --batch 128 > gpurun_out/${tag}_ncu_metrics.log 2>&1
echo "metrics exit=$?"
timeout 1500 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:'conv_igemm' -s $skip -c $cnt -o gpurun_out/${tag}_prof -f python tools/profile_step.py --batch 128 > gpurun_out/${tag}_ncu_full.log 2>&1
echo "full capture exit=$?"; tail -n 3 gpurun_out/${tag}_ncu_full.log
du -sh gpurun_out; ls -la gpurun_out
