mkdir -p gpurun_out
( nvidia-smi -L
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/peer_check.py 2>&1 | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/ddp_check.py 20 2>&1 | tail -4
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 bench.py --gpus 2 --steps 5 --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/b2.log 2> gpurun_out/b2.err; echo "bench2 rc=$?"; tail -3 gpurun_out/b2.err
python -c "
import json; d=json.loads(open('gpurun_out/b2.log').read().strip().splitlines()[-1]); s=d['summary']; print(json.dumps(s['train_strong'])); print(json.dumps(s['train_weak'])); print(json.dumps(s['ddp_parity'])); print(json.dumps(s['sampling']))"
echo "--- NCCL BatchNorm exchange"
MSTK_PEER_BN=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29537 bench.py --gpus 2 --steps 5 --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/b2n.log 2> gpurun_out/b2n.err; echo "bench2 rc=$?"; tail -3 gpurun_out/b2n.err
python -c "
import json; d=json.loads(open('gpurun_out/b2n.log').read().strip().splitlines()[-1]); s=d['summary']; print(json.dumps(s['train_strong'])); print(json.dumps(s['train_weak'])); print(json.dumps(s['ddp_parity']))"
) > gpurun_out/peer.log 2>&1
cat gpurun_out/peer.log | cut -c 1-500
