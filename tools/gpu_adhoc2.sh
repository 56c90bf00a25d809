mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/ddp_check.py 30 > gpurun_out/ddp_check.log 2>&1; echo "ddp_check rc=$?"; grep -E "ddp_check world|MstkError" gpurun_out/ddp_check.log | head -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/b_2gpu.log 2>gpurun_out/b_2gpu.err; echo "bench2 rc=$?"; grep -E "MstkError|Error" gpurun_out/b_2gpu.err | head -3
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/b_2gpu.log').read().strip().splitlines()[-1])
    print("n_gpus", d['n_gpus'], "steps/s", d['value'], "e2e", d['e2e']['value'])
    print("train", d['train']['value'], d['train']['ms_per_step'], d['train']['e2e'])
except Exception as e: print("no bench", e)
PY
