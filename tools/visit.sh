mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_backward_ops.py tests/test_gpu_ops_fp32.py tests/test_gpu_train.py tests/test_gpu_train_autoencoder.py tests/test_gpu_baseline_sizes.py -q -m gpu -x --timeout 600 2>&1 | tail -2
bash tools/bench_train.sh 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"bn_" -c 26 python tools/profile_train.py 256 2 2>&1 | grep -i "gpu__time_duration\|bn_" | paste - - | awk '{print $1, $NF}' | tail -13 ) > gpurun_out/bn.log 2>&1
cat gpurun_out/bn.log | cut -c 1-300
