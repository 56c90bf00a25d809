mkdir -p gpurun_out
MSTK_TC_DIRECT=1 timeout 600 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_end_to_end.py -x -q 2>&1 | tail -2
for dflag in 1 0; do
MSTK_TC_DIRECT=$dflag timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_adhoc.json 2> gpurun_out/bench_adhoc.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_adhoc.json').read().strip().splitlines()[-1])
t=d.get('train',{})
print('direct=$dflag sampling', d['value'], 'e2e', d['e2e']['value'], '| train', t.get('value'), t.get('ms_per_step'), t['roofline']['frac'], [round(l['tflops']) for l in t['roofline']['layers']])
PY
done
