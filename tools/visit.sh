mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --no-train --no-vae --no-mel --no-gpu-baseline --no-cpu-baseline > gpurun_out/bk.log 2> gpurun_out/bk.err; echo rc=$?; tail -3 gpurun_out/bk.err | cut -c 1-200
python -c "
import json; d=json.loads(open('gpurun_out/bk.log').read().strip().splitlines()[-1]); print(json.dumps(d['batch_sweep'])); print(d['summary']['sampling'])"
