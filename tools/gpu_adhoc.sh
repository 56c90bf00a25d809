mkdir -p gpurun_out
timeout 600 python tools/tc_tune.py 64 > gpurun_out/tune64.log 2>&1; echo "tune rc=$?"; cat gpurun_out/tune64.log
