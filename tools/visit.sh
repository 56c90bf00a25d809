mkdir -p gpurun_out
( s=$(date +%s)
timeout 1200 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/full.log 2>&1; echo "pytest rc=$? $(( $(date +%s) - s ))s"; tail -2 gpurun_out/full.log | cut -c 1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
s=$(date +%s)
timeout 1500 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$? $(( $(date +%s) - s ))s"
python -c "
import json; d=json.loads(open('gpurun_out/bench.log').read().strip().splitlines()[-1]); print(json.dumps(d['summary']['sampling'])); print(json.dumps(d['summary']['train_strong'])); print(json.dumps(d['summary']['vae'])); e=d['e2e']; print(e['value'], e['method'], [round(v,3) for v in e['runs_ms_per_step']], e['serial_value']); print(json.dumps(d['train']['e2e'])[:500]); print(json.dumps(d['clocks'])); print(d['gpu_launches'], d['value'])"
) > gpurun_out/final.log 2>&1
cat gpurun_out/final.log | cut -c 1-1500
