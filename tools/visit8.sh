mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
( echo "gpus $N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/b8.log 2> gpurun_out/b8.err; echo "bench$N rc=$?"; tail -3 gpurun_out/b8.err | grep -v "OMP_NUM\|\*\*\*\*"
python -c "
import json; d=json.loads(open('gpurun_out/b8.log').read().strip().splitlines()[-1]); s=d['summary']; print(json.dumps(s['train_strong'])); print(json.dumps(s['train_weak'])); print(json.dumps(s['ddp_parity'])); print(json.dumps(s['sampling'])); print(json.dumps(d['e2e'])[:400])"
) > gpurun_out/peer8.log 2>&1
cat gpurun_out/peer8.log | cut -c 1-700
