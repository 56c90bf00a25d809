mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_train.py -q -m gpu --timeout 600 -s > gpurun_out/t_train.log 2>&1; echo "train tests rc=$?"; tail -n 40 gpurun_out/t_train.log
