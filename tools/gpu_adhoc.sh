mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_wgrad_tc.py tests/test_gpu_train.py -x -q 2>&1 | tail -2
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_adhoc.json 2> gpurun_out/bench_adhoc.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_adhoc.json').read().strip().splitlines()[-1])
t=d.get('train',{})
print('sampling', d['value'], '| train', t.get('value'), t.get('ms_per_step'), t['roofline']['frac'])
PY
