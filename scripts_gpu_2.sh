#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_model.py -q -m gpu --timeout 300 -k "bf16 or full_size" > gpurun_out/t_model_bf16.log 2>&1; echo "model_bf16 rc=$?" | tee gpurun_out/rc.txt
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/rc.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/b_short.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?" | tee -a gpurun_out/rc.txt
tail -n 3 gpurun_out/t_model_bf16.log; cat gpurun_out/smoke.log | tail -n 4; cat gpurun_out/bench.log | cut -c 1-1500; tail -n 5 gpurun_out/bench.err
