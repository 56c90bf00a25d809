mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_loss_fuzz.py -q -m gpu --timeout 600 -p no:cacheprovider > gpurun_out/fuzz.log 2>&1
grep -E "^FAILED|^ERROR|passed|failed" gpurun_out/fuzz.log | cut -c 1-330 | head -50
grep -n -E "^E  " gpurun_out/fuzz.log | head -40 | cut -c 1-300
