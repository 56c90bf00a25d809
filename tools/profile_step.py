#!/usr/bin/env python
"""Runs a few EAGER batch-64 bf16 UNet+DDIM steps (no CUDA graph) so that ncu can pick
individual launches:  ncu -k regex:conv_tc_kernel -s <skip> -c <n> python tools/profile_step.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import music_style_transfer_ldm_b200 as M
from oracle import ldm_oracle as O

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16"
dev = "cuda"
m = M.LDM(32, pretrained_path=None)
m.load_state_dict(O.make_state_dict(1234))
m.to(dev).eval().set_precision(prec)
m.use_cuda_graphs = False
g = torch.Generator().manual_seed(1)
z = torch.randn((64, 32, 16, 16), generator=g).to(dev)
st = {"s5": torch.rand((64, 256, 4, 4), generator=g).to(dev), "s6": torch.rand((64, 512, 2, 2), generator=g).to(dev)}
with torch.no_grad():
    x, _ = m.content_style_ddim_sample(z, st, timesteps=steps + 1, eta=0.0)
torch.cuda.synchronize()
print("ok", float(x.abs().mean()))
