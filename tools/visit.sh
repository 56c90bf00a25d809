mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_model.py -q -m gpu -k transfer_batches --timeout 300 -p no:cacheprovider 2>&1 | tail -2
run() { timeout 600 python bench.py --steps 20 --no-train --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bk.log 2> gpurun_out/bk.err
python -c "
import json; d=json.loads(open('gpurun_out/bk.log').read().strip().splitlines()[-1]); e=d['e2e']; print(round(d['ms_per_step'],3), round(e['ms_per_step'],3), [round(v,3) for v in e['runs_ms_per_step']], round(e['serial_ms_per_step'],3))"; }
run; run; run; run
