mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_model.py tests/test_gpu_program.py -q -m gpu --timeout 600 -x > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/t_gpu.log
for k in 1 0; do
MSTK_KROT=$k timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-train > gpurun_out/b_krot$k.log 2>gpurun_out/b_short.err; echo "bench rc=$?"; tail -n 5 gpurun_out/b_short.err
python - <<PY
import json
d=json.loads(open('gpurun_out/b_krot$k.log').read().strip().splitlines()[-1])
print("krot=$k steps/s", d['value'], "e2e", d['e2e']['value'], "us/step", d['unet_step_us'], "launches", d['gpu_launches'])
print(" ".join(f"{r['us']:.1f}" for r in d['layers']))
PY
done
