mkdir -p gpurun_out
( s=$(date +%s)
timeout 1200 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/full.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - s ))s"; tail -4 gpurun_out/full.log | cut -c 1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log | cut -c 1-250
s=$(date +%s)
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "ref rc=$? $(( $(date +%s) - s ))s"; cut -c 1-300 gpurun_out/bench_ref.log
s=$(date +%s)
timeout 1500 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$? $(( $(date +%s) - s ))s"
python -c "
import json; d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1]); print(json.dumps(d['summary'])[:3000]); print(json.dumps(d['e2e'])[:900]); print(json.dumps(d['train']['e2e'])[:700]); print(json.dumps(d['roofline'])[:300]); print(json.dumps(d['clocks'])[:300]); print(d['gpu_launches'], d['value'], d['steps'], d['warmup'])"
) > gpurun_out/final.log 2>&1
cat gpurun_out/final.log | cut -c 1-3300
