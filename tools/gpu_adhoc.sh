mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --timeout 600 -x > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"; tail -n 12 gpurun_out/t_gpu.log
timeout 300 python tools/profile_train.py 256 3 > gpurun_out/pt_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/train_launches.csv python tools/profile_train.py 256 3 > gpurun_out/pt_ncu.log 2>&1; echo "ncu rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b_short.log 2>&1; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/b_short.log').read().strip().splitlines()[-1])
print("steps/s", d['value'], "train", d['train']['value'], d['train']['ms_per_step'])
PY
