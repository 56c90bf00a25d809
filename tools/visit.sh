mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_autoencoder.py tests/test_gpu_autograd.py -q -m gpu -x --timeout 600 2>&1 | tail -2
bash tools/bench_train.sh 2>&1 | tail -2 ) > gpurun_out/wg3.log 2>&1
cat gpurun_out/wg3.log | cut -c 1-300
