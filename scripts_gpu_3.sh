#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/profile_step.py 4 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv_tc_kernel|attention_kernel" -s 36 -c 17 -o gpurun_out/prof_r1_step python tools/profile_step.py 4 > gpurun_out/prof_ncu.log 2>&1
echo "rc=$?"; tail -n 3 gpurun_out/prof_ncu.log
