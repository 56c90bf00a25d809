mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_conv_edge.py -q -m gpu -x --timeout 300 2>&1 | tail -8
timeout 1500 python -m pytest tests/test_gpu_model.py tests/test_gpu_baseline_sizes.py tests/test_gpu_train.py tests/test_gpu_train_autoencoder.py tests/test_gpu_autograd.py tests/test_gpu_end_to_end.py -q -m gpu -x --timeout 600 2>&1 | tail -2
bash tools/bench_train.sh 2>&1 | tail -2
timeout 600 python bench.py --steps 3 --no-train --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bv.log 2> gpurun_out/bv.err; python -c "
import json; d=json.loads(open('gpurun_out/bv.log').read().strip().splitlines()[-1]); print(json.dumps(d['summary']['vae'])); print(json.dumps(d['summary']['sampling']))" ) > gpurun_out/c1.log 2>&1
cat gpurun_out/c1.log | cut -c 1-400
