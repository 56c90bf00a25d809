#!/bin/bash
# Short GPU-box visit: parity tests, smoke, bench (both arms), ncu launch list of the bench command.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
S=$(date +%s)
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/t_gpu.log 2>&1; echo "pytest_gpu rc=$? t=$(( $(date +%s)-S ))" | tee gpurun_out/rc.txt
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$? t=$(( $(date +%s)-S ))" | tee -a gpurun_out/rc.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$? t=$(( $(date +%s)-S ))" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$? t=$(( $(date +%s)-S ))" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/b_short.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo "ncu_list rc=$? t=$(( $(date +%s)-S ))" | tee -a gpurun_out/rc.txt
tail -n 5 gpurun_out/t_gpu.log; tail -n 4 gpurun_out/smoke.log; cut -c 1-1800 gpurun_out/bench.log; tail -n 5 gpurun_out/bench.err; cut -c 1-600 gpurun_out/bench_ref.log
du -sh gpurun_out | tee -a gpurun_out/rc.txt
