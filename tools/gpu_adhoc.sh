mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_conv_tc.py -q -m gpu --timeout 120 -x > gpurun_out/t_tc.log 2>&1; echo "tc rc=$?"
tail -n 25 gpurun_out/t_tc.log
timeout 300 python tools/tc_timeline.py > gpurun_out/timeline.log 2>&1; echo "timeline rc=$?"; cat gpurun_out/timeline.log
timeout 900 python -m pytest tests/test_gpu_model.py -q -m gpu --timeout 300 -x > gpurun_out/t_model.log 2>&1; echo "model rc=$?"; tail -n 15 gpurun_out/t_model.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b_short.log 2>gpurun_out/b_short.err; echo "bench rc=$?"; tail -n 3 gpurun_out/b_short.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/b_short.log').read().strip().splitlines()[-1])
    print("steps/s", d['value'], "e2e", d['e2e']['value'], "unet_step_us", d['unet_step_us'], "roof", d['roofline']['achieved'])
    for r in d['layers']: print(f"{r['layer']:45s} {r['us']:8.1f} us {r['tflops']:8.2f} TF/s")
except Exception as e: print("no bench", e)
PY
