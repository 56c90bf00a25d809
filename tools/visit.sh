mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_train.py tests/test_gpu_autograd.py tests/test_gpu_baseline_sizes.py tests/test_gpu_feature_loss.py -q -m gpu -x --timeout 600 2>&1 | tail -2
bash tools/bench_train.sh 2>&1 ) > gpurun_out/ws.log 2>&1
cat gpurun_out/ws.log | cut -c 1-400
