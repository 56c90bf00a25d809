mkdir -p gpurun_out
timeout 300 python tools/tc_timeline.py > gpurun_out/timeline.log 2>&1; echo "timeline rc=$?"
timeout 600 python -m pytest tests/test_gpu_mel.py -q -m gpu --timeout 300 > gpurun_out/t_mel.log 2>&1; echo "mel rc=$?"
cat gpurun_out/timeline.log; tail -n 30 gpurun_out/t_mel.log
