mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_ddp.py tests/test_dataset.py -q -m gpu --timeout 500 -p no:cacheprovider 2>&1 | tail -3
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29545 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/b2full.log 2> gpurun_out/b2full.err; echo "bench2 full rc=$?"; tail -3 gpurun_out/b2full.err | grep -v "OMP\|\*\*\*"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29547 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/b2ref.log 2> gpurun_out/b2ref.err; echo "bench2 ref rc=$?"; cut -c 1-200 gpurun_out/b2ref.log
python -c "
import json; d=json.loads(open('gpurun_out/b2full.log').read().strip().splitlines()[-1]); s=d['summary']; print(json.dumps(s)[:1800])"
) > gpurun_out/full2.log 2>&1
cat gpurun_out/full2.log | cut -c 1-1900
