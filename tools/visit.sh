mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_gpu_conv_halo.py tests/test_gpu_train.py tests/test_gpu_baseline_sizes.py tests/test_gpu_train_autoencoder.py -q -m gpu -x --timeout 900 2>&1 | tail -2
timeout 600 python bench.py --steps 3 --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bv.log 2> gpurun_out/bv.err; tail -2 gpurun_out/bv.err
python -c "
import json; d=json.loads(open('gpurun_out/bv.log').read().strip().splitlines()[-1]); print(json.dumps(d['summary']['vae'])); print(json.dumps(d['summary']['train_strong'])); print(d['summary']['autoencoder_samples_per_s'])" ) > gpurun_out/vae.log 2>&1
cat gpurun_out/vae.log | cut -c 1-600
