#!/bin/bash
# full ncu capture of chosen conv launches of one forward: bash tools/gpu_capture.sh <tag> <skip> <count> [regex]
tag=$1; skip=$2; cnt=$3; rx=${4:-conv_igemm}
mkdir -p gpurun_out
timeout 300 python tools/profile_step.py --batch 128 > gpurun_out/${tag}_plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:$rx -s $skip -c $cnt -o gpurun_out/${tag}_prof -f python tools/profile_step.py --batch 128 > gpurun_out/${tag}_ncu_full.log 2>&1
echo "capture exit=$?"; tail -n 2 gpurun_out/${tag}_ncu_full.log
