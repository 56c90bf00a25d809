mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_train_autoencoder.py tests/test_gpu_ddp.py -q -m gpu --timeout 600 -p no:cacheprovider 2>&1 | tail -3
timeout 900 python bench.py --steps 6 --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bv.log 2> gpurun_out/bv.err; echo "bench rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bv.log').read().strip().splitlines()[-1]); print(json.dumps(d['train']['e2e'])); print(json.dumps(d['summary']['train'])); print(json.dumps(d['summary']['sampling']))" ) > gpurun_out/c1.log 2>&1
cat gpurun_out/c1.log | cut -c 1-900
