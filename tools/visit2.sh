mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_ddp.py -q -m gpu --timeout 800 2>&1 | tail -4
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus 2 --steps 5 --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/b2.log 2> gpurun_out/b2.err; echo "bench2 rc=$?"; tail -3 gpurun_out/b2.err | grep -v "OMP\|\*\*\*"
python -c "
import json; d=json.loads(open('gpurun_out/b2.log').read().strip().splitlines()[-1]); s=d['summary']; print(json.dumps(s['train_strong'])); print(json.dumps(s['train_weak'])); print(json.dumps(s['ddp_parity'])); print(json.dumps(s['sampling']))"
) > gpurun_out/ddp2.log 2>&1
cat gpurun_out/ddp2.log | cut -c 1-500
