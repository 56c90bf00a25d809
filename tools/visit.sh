mkdir -p gpurun_out
timeout 900 python tools/capture_probe.py tests/test_gpu_model.py tests/test_gpu_baseline_sizes.py tests/test_gpu_train.py tests/test_gpu_train_autoencoder.py tests/test_gpu_autograd.py tests/test_gpu_end_to_end.py -q -m gpu -x --timeout 600 -p no:cacheprovider > gpurun_out/probe.log 2>&1
grep -n -A45 "####" gpurun_out/probe.log | cut -c 1-220 | head -120
tail -3 gpurun_out/probe.log | cut -c 1-200
