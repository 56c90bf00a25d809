mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -n 6 gpurun_out/t_gpu.log
timeout 900 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/b_short.log 2>gpurun_out/b_short.err; echo "bench rc=$?"; tail -n 5 gpurun_out/b_short.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/b_short.log').read().strip().splitlines()[-1])
    print("steps/s", d['value'], "e2e", d['e2e']['value'], "unet_step_us", d['unet_step_us'], "roof", d['roofline']['achieved'])
    print("train", json.dumps(d['train'])[:900])
except Exception as e: print("no bench", e)
PY
