mkdir -p gpurun_out
for c in 1800 4000 6600 10000; do
MSTK_TC_SPLIT_COST=$c MSTK_TC_TRACE=1 timeout 600 python bench.py --no-cpu-baseline --no-train > gpurun_out/bench_adhoc.json 2> gpurun_out/bench_adhoc.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_adhoc.json').read().strip().splitlines()[-1])
print('split cost $c: sampling', d['value'], 'e2e', d['e2e']['value'])
PY
grep "N64 4x4 Cin256 Cout512\|N64 2x2 Cin512 Cout512 k1\|N64 2x2 Cin512 Cout256\|N64 4x4 Cin256 Cout128" gpurun_out/bench_adhoc.err | sort | uniq -c | sort -rn | head -4 | cut -c 1-150
done
