mkdir -p gpurun_out
: > gpurun_out/flaky.log
for i in 1 2 3; do
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_baseline_sizes.py tests/test_gpu_train.py tests/test_gpu_train_autoencoder.py tests/test_gpu_autograd.py tests/test_gpu_end_to_end.py -q -m gpu --timeout 600 > gpurun_out/f$i.log 2>&1
tail -1 gpurun_out/f$i.log >> gpurun_out/flaky.log
grep -n -E "^E  |^FAILED" gpurun_out/f$i.log | cut -c 1-250 | head -20 >> gpurun_out/flaky.log
done
timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/full.log 2>&1; tail -4 gpurun_out/full.log | cut -c 1-200 >> gpurun_out/flaky.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 >> gpurun_out/flaky.log
cat gpurun_out/flaky.log
