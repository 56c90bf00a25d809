# 2-GPU check of the data-parallel path at HEAD: DDP parity tests, ddp_check vs the oracle, bench.py at N = 2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ddp.py -q -m gpu 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/ddp_check.py 20 2>&1 | tail -4
n=2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 6 --warmup 3 > gpurun_out/b_${n}gpu.log 2>gpurun_out/b_${n}gpu.err; echo "bench$n rc=$?"; grep -E "MstkError|Error" gpurun_out/b_${n}gpu.err | head -3
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/b_${n}gpu.log').read().strip().splitlines()[-1])
    print("n_gpus", d['n_gpus'], "steps/s", d['value'], "e2e", d['e2e']['value'], "traffic", d['roofline']['traffic'])
    print("train", d['train']['value'], d['train']['ms_per_step'], d['train']['e2e']['value'])
    w=d.get('train_weak')
    if w: print("train_weak", w['value'], w['ms_per_step'], w['global_batch'])
except Exception as e: print("no bench", e)
PY
