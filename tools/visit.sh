mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_autoencoder.py tests/test_gpu_autograd.py tests/test_gpu_baseline_sizes.py tests/test_gpu_model.py -q -m gpu -x --timeout 900 2>&1 | tail -6
bash tools/bench_train.sh 2>&1 ) > gpurun_out/pack.log 2>&1
cat gpurun_out/pack.log | cut -c 1-400
