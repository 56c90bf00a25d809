mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_train.py -q -m gpu --timeout 600 -p no:cacheprovider 2>&1 | tail -3
timeout 900 python bench.py --steps 6 --no-train --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bv.log 2> gpurun_out/bv.err; echo "bench rc=$?"; tail -3 gpurun_out/bv.err; python -c "
import json; d=json.loads(open('gpurun_out/bv.log').read().strip().splitlines()[-1]); print(json.dumps(d['e2e'])); print(json.dumps(d['summary']['sampling']))" ) > gpurun_out/c1.log 2>&1
cat gpurun_out/c1.log | cut -c 1-900
