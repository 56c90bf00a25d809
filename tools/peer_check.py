#!/usr/bin/env python
"""2..8 GPUs under torchrun: the peer-memory all-reduce (csrc/peer.cu) against NCCL on the same fp64 vectors (eager and
replayed from a CUDA graph), and the latency of both for a BatchNorm statistics vector (256 doubles).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/peer_check.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from music_style_transfer_ldm_b200.train import PeerExchange

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
px = PeerExchange(None, world, dev)
g = torch.Generator(device=dev).manual_seed(100 + rank)
ok = True
# (1) eager: 5 "steps" of 9 exchanges of different lengths
for step in range(5):
    px.begin_step()
    for n in (128, 256, 64, 256, 128, 128, 256, 2, 512):
        t = torch.randn(n, dtype=torch.float64, device=dev, generator=g)
        ref = t.clone()
        dist.all_reduce(ref)
        px.all_reduce_(t)
        # same rank-ordered sum on every rank; NCCL's order may differ in the last bits
        ok &= bool(torch.allclose(t, ref, rtol=1e-13, atol=1e-13))
        chk = t.clone(); lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        ok &= bool(torch.equal(lo, hi))                      # bit-identical on every rank
# (2) captured: the same slots replayed 20 times
bufs = [torch.zeros(256, dtype=torch.float64, device=dev) for _ in range(9)]
src = [torch.randn(256, dtype=torch.float64, device=dev, generator=g) for _ in range(9)]
torch.cuda.synchronize()
graph = torch.cuda.CUDAGraph()
px.begin_step()
with torch.cuda.graph(graph):
    for b, s_ in zip(bufs, src):
        b.copy_(s_)
        px.all_reduce_(b)
for _ in range(20):
    graph.replay()
torch.cuda.synchronize()
for b, s_ in zip(bufs, src):
    ref = s_.clone(); dist.all_reduce(ref)
    ok &= bool(torch.allclose(b, ref, rtol=1e-13, atol=1e-13))
# (3) latency: 9 dependent exchanges per replay, as in one training step
def time_graph(fn, reps=50):
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        fn()
    for _ in range(5):
        gr.replay()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        gr.replay()
    e1.record(); torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / reps / 9

def peer9():
    px.begin_step()
    for b in bufs:
        px.all_reduce_(b)

def nccl9():
    for b in bufs:
        dist.all_reduce(b)

us_peer = time_graph(peer9)
us_nccl = time_graph(nccl9)
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"peer_check world={world}: peer-memory all-reduce {us_peer:.2f} us per exchange, NCCL {us_nccl:.2f} us per exchange (256 doubles, "
          f"9 dependent exchanges per graph replay) -> {'PASS' if int(flag) else 'FAIL'}")
px.close()
dist.destroy_process_group()
sys.exit(0 if int(flag) else 1)
