#!/bin/bash
# One GPU-box visit, parameterised:  tools/gpu_check.sh [tests] [smoke] [bench] [ref] [ncu] [pytest-args...]
# Everything lands in gpurun_out/; numbers printed under ncu are never bench values.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
: > gpurun_out/rc.txt
for what in "$@"; do
  case "$what" in
    tests) timeout 2400 python -m pytest tests -q -m gpu --timeout 900 -x > gpurun_out/t_gpu.log 2>&1; echo "pytest_gpu rc=$?" | tee -a gpurun_out/rc.txt; tail -n 60 gpurun_out/t_gpu.log ;;
    newtests) timeout 1800 python -m pytest tests/test_gpu_baseline_sizes.py -q -m gpu --timeout 900 > gpurun_out/t_new.log 2>&1; echo "pytest_new rc=$?" | tee -a gpurun_out/rc.txt; tail -n 70 gpurun_out/t_new.log ;;
    smoke) timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/rc.txt; tail -n 4 gpurun_out/smoke.log ;;
    bench) timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/rc.txt
           tail -n 5 gpurun_out/bench.err; python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench.log").read().strip().splitlines()[-1])
    print(json.dumps(d["summary"], indent=1)); print(json.dumps(d.get("roofline"))[:1500])
    for r in d["layers"]: print(f'{r["us"]:7.2f} us  {r["tflops"]:7.1f} TF  {r["layer"]}')
except Exception as e:
    print("bench parse failed:", e)
PY
           ;;
    ref) timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$?" | tee -a gpurun_out/rc.txt; cut -c 1-400 gpurun_out/bench_ref.log ;;
    ncu) timeout 300 python tools/profile_step.py 4 > gpurun_out/prof_plain.log 2>&1 && \
         timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/step_launches.csv python tools/profile_step.py 4 > gpurun_out/ncu.log 2>&1; echo "ncu_list rc=$?" | tee -a gpurun_out/rc.txt
         timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv_tc_kernel|attn" -s 30 -c 14 -o gpurun_out/prof_step python tools/profile_step.py 4 > gpurun_out/prof_ncu.log 2>&1; echo "ncu_full rc=$?" | tee -a gpurun_out/rc.txt
         [ -f gpurun_out/prof_step.ncu-rep ] && ncu -i gpurun_out/prof_step.ncu-rep --page raw --csv > gpurun_out/prof_step_raw.csv 2>/dev/null ;;
    *) echo "unknown step $what" ;;
  esac
done
du -sh gpurun_out | tee -a gpurun_out/rc.txt
