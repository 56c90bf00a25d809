mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --timeout 600 -x > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/t_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-train > gpurun_out/b_short.log 2>gpurun_out/b_short.err; echo "bench rc=$?"; tail -n 5 gpurun_out/b_short.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/b_short.log').read().strip().splitlines()[-1])
print("steps/s", d['value'], "e2e", d['e2e']['value'], "us/step", d['unet_step_us'])
PY
