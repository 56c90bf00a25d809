mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_adhoc.json 2> gpurun_out/bench_adhoc.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_adhoc.json').read().strip().splitlines()[-1])
print('sampling', d['value'], 'e2e', d['e2e']['value'], 'roof', d['roofline']['frac'])
t=d.get('train',{})
print('train', t.get('value'), t.get('ms_per_step'), t['roofline']['frac'], [round(l['tflops']) for l in t['roofline']['layers']])
PY
