#!/bin/bash
# mel / config-5 leg only
mkdir -p gpurun_out
python bench.py --steps 3 --no-train --no-vae --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bm.log 2>gpurun_out/bm.err; tail -3 gpurun_out/bm.err
python -c "
import json; d=json.loads(open('gpurun_out/bm.log').read().strip().splitlines()[-1]); print(json.dumps(d['summary']['config5'])); print(json.dumps(d['config5']['mel']))"
