mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_conv_halo.py tests/test_gpu_conv_tc.py tests/test_gpu_feature_loss.py tests/test_gpu_train.py tests/test_gpu_fp32_tc.py -q -m gpu -x --timeout 600 2>&1 | tail -8
bash tools/bench_train.sh 2>&1
echo "--- halo off"; MSTK_CONV_HALO=0 bash tools/bench_train.sh 2>&1 ) > gpurun_out/halo_train.log 2>&1
cat gpurun_out/halo_train.log | cut -c 1-400
