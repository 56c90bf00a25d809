mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_model.py -q -m gpu --timeout 600 -s > gpurun_out/t_train.log 2>&1; echo "tests rc=$?"; grep -E "worst|passed|failed|Error" gpurun_out/t_train.log | head
timeout 300 python tools/profile_train.py 256 3 > gpurun_out/pt_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/train_launches.csv python tools/profile_train.py 256 3 > gpurun_out/pt_ncu.log 2>&1; echo "ncu rc=$?"
