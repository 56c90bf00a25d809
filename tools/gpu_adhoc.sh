mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_backward_ops.py -q -m gpu --timeout 600 -x > gpurun_out/t_train.log 2>&1; echo "tests rc=$?"; tail -n 15 gpurun_out/t_train.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b_short.log 2>gpurun_out/b_short.err; echo "bench rc=$?"; tail -n 5 gpurun_out/b_short.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/b_short.log').read().strip().splitlines()[-1])
t=d['train']
print("steps/s", d['value'], "train", t['value'], t['ms_per_step'], "e2e", t['e2e']['value'], 'launches', t['gpu_launches_per_step'], t['last_losses'])
PY
