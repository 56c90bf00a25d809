mkdir -p gpurun_out
run() { timeout 600 env $1 python bench.py --steps 20 --no-train --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bk.log 2> gpurun_out/bk.err
python -c "
import json; d=json.loads(open('gpurun_out/bk.log').read().strip().splitlines()[-1]); e=d['e2e']; print('$1', round(d['ms_per_step'],3), [round(v,3) for v in e['runs_ms_per_step']], round(e['serial_ms_per_step'],3))"; }
for i in 1 2 3 4; do run A=1; run MSTK_BENCH_NOSAMPLER=1; done
