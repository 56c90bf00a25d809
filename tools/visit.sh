mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_conv_halo.py -q -m gpu -x --timeout 300 2>&1 | tail -12
timeout 1200 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_autoencoder.py tests/test_gpu_autograd.py tests/test_gpu_baseline_sizes.py tests/test_gpu_model.py -q -m gpu -x --timeout 600 2>&1 | tail -2
bash tools/bench_train.sh 2>&1 ) > gpurun_out/k3.log 2>&1
cat gpurun_out/k3.log | cut -c 1-400
