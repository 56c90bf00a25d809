#!/bin/bash
# sampling-only bench (no train / VAE / mel / baselines): prints the summary and the per-layer table
mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 --no-train --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline "$@" > gpurun_out/bq.log 2> gpurun_out/bq.err; echo "bench rc=$?"
tail -n 5 gpurun_out/bq.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bq.log").read().strip().splitlines()[-1])
print(json.dumps(d["summary"]["sampling"])); r = d["roofline"]
print({k: r[k] for k in ("frac", "chained_step_us", "sum_isolated_us_all_kernels", "launches_per_unet_step_all_kernels")})
for r in d["layers"]: print(f'{r["us"]:7.2f} us  {r["tflops"]:7.1f} TF  {r["layer"]}')
PY
