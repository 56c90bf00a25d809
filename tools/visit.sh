mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv_halo.py tests/test_gpu_conv_tcp.py tests/test_gpu_conv_tc.py -q -m gpu --timeout 300 2>&1 | tail -3 > gpurun_out/s2.log; cat gpurun_out/s2.log
