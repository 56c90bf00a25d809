mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_backward_ops.py -q -m gpu --timeout 300 > gpurun_out/t_bwd.log 2>&1; echo "bwd tests rc=$?"; tail -n 40 gpurun_out/t_bwd.log
