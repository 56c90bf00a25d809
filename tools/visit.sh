mkdir -p gpurun_out
for i in 1 2 3; do
timeout 600 python bench.py --steps 20 --no-train --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bk.log 2> gpurun_out/bk.err; echo rc=$?
python -c "
import json; d=json.loads(open('gpurun_out/bk.log').read().strip().splitlines()[-1]); e=d['e2e']; print([round(v,3) for v in e['runs_ms_per_step']], round(e['serial_ms_per_step'],3), e['method'])"
done
