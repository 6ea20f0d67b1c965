#!/bin/bash
# full ncu capture of two fused-MAT conv launches (25x25 C=512 and 100x100 C=128) in high precision + graph test
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_generator_gpu.py -m gpu -q -x -k "graph" -s > gpurun_out/r02e_pytest.log 2>&1; echo "pytest exit=$?"; tail -n 5 gpurun_out/r02e_pytest.log
timeout 300 python tools/profile_step.py --batch 128 --precision high > gpurun_out/r02e_plain.log 2>&1
for s in 6 22; do
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'conv_igemm' -s $s -c 1 \
   -o gpurun_out/r02e_fused_$s -f python tools/profile_step.py --batch 128 --precision high > gpurun_out/r02e_ncu_$s.log 2>&1
echo "capture $s exit=$?"
done
ls -la gpurun_out/*.ncu-rep
