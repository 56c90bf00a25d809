# scaling check on one box: bench.py at N = 4 and 8 (one process per GPU, NCCL)
mkdir -p gpurun_out
for n in 4 8; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 6 --warmup 3 > gpurun_out/b_${n}gpu.log 2>gpurun_out/b_${n}gpu.err; echo "bench$n rc=$?"; grep -E "MstkError|Error" gpurun_out/b_${n}gpu.err | head -3
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/b_${n}gpu.log').read().strip().splitlines()[-1])
    print("n_gpus", d['n_gpus'], "steps/s", d['value'], "e2e", d['e2e']['value'])
    print("train", d['train']['value'], d['train']['ms_per_step'], d['train']['e2e']['value'])
    w=d.get('train_weak')
    if w: print("train_weak", w['value'], w['ms_per_step'], w['global_batch'])
except Exception as e: print("no bench", e)
PY
done
