mkdir -p gpurun_out
( timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 -p no:cacheprovider > gpurun_out/full.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/full.log | cut -c 1-150
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1]); s=d['summary']; print(json.dumps(s['sampling'])); print(json.dumps(s['train_strong'])); print(d['e2e']['value'], d['e2e']['method'])" ) > gpurun_out/final.log 2>&1
cat gpurun_out/final.log | cut -c 1-900
