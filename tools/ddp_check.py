#!/usr/bin/env python
"""Data-parallel parity of the training step, run under torchrun on N GPUs of one node:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/ddp_check.py [steps]

The GLOBAL batch of tests/golden/make_golden_curve.py (4 samples, pre-drawn t / noise) is sharded over the ranks; with the
BatchNorm statistics and the gradients all-reduced, the per-step losses must follow the single-process oracle curve
(tests/golden/train_curve_100.npz) within 1 %, and every rank must hold identical parameters afterwards."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests", "golden")]
import numpy as np
import torch
import torch.distributed as dist

import cases
import make_golden_curve as GC
import music_style_transfer_ldm_b200 as M
from music_style_transfer_ldm_b200.train import LDMTrainer
from oracle import ldm_oracle as O

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
assert GC.B % world == 0
per = GC.B // world
m = M.LDM(latent_dim=32, pretrained_path=None)
m.load_state_dict(O.make_state_dict(cases.WEIGHT_SEED), strict=True)
m._freeze_encoder()
m.to(dev).set_precision("bf16")
tr = LDMTrainer(m, None, dev, lr=GC.LR, style_loss_weight=GC.W, process_group=None, world_size=world)
m.train()
gold = cases.load("train_curve_100")["losses"]
worst = 0.0
for i in range(steps):
    x, s, t, eps = GC.step_inputs(i)
    sl = slice(rank * per, (rank + 1) * per)
    out = tr.train_step_device(x[sl].to(dev), s[sl].to(dev), t=t[sl].to(dev), noise=eps[sl].to(dev))
    got = np.array([float(out[k]) for k in ("compression_loss", "denoisinsg_loss", "style_loss", "total_loss")])
    worst = max(worst, float((np.abs(got - gold[i]) / np.abs(gold[i])).max()))
chk = tr.arena.p.double().sum().reshape(1)
lo, hi = chk.clone(), chk.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
tr.close()          # the captured step holds NCCL kernels: it must go before the process group does
ok = worst <= 1e-2 and float(hi - lo) == 0.0
if rank == 0:
    print(f"ddp_check world={world} steps={steps}: worst loss deviation vs single-process oracle {worst:.3e}; "
          f"replica checksum spread {float(hi - lo):.3e} -> {'PASS' if ok else 'FAIL'}")
dist.destroy_process_group()
sys.exit(0 if ok else 1)
