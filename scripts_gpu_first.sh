#!/bin/bash
# first GPU contact: fp32 kernels, then tcgen05 kernel (own process: a trap poisons the context), then models
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_ops_fp32.py -q -m gpu -x --timeout 300 > gpurun_out/t_fp32.log 2>&1; echo "fp32 rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_gpu_conv_tc.py -q -m gpu --timeout 120 > gpurun_out/t_tc.log 2>&1; echo "tc rc=$?" | tee -a gpurun_out/rc.txt
timeout 900 python -m pytest tests/test_gpu_model.py -q -m gpu --timeout 300 -k "fp32" > gpurun_out/t_model_fp32.log 2>&1; echo "model_fp32 rc=$?" | tee -a gpurun_out/rc.txt
timeout 900 python -m pytest tests/test_gpu_model.py -q -m gpu --timeout 300 -k "bf16 or full_size" > gpurun_out/t_model_bf16.log 2>&1; echo "model_bf16 rc=$?" | tee -a gpurun_out/rc.txt
tail -5 gpurun_out/t_fp32.log gpurun_out/t_tc.log gpurun_out/t_model_fp32.log gpurun_out/t_model_bf16.log
