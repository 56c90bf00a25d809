mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu --timeout 900 > gpurun_out/t_gpu.log 2>&1; echo "pytest_gpu rc=$?" | tee gpurun_out/rc.txt
tail -n 6 gpurun_out/t_gpu.log
