#!/usr/bin/env python
"""Runs a few bf16 training steps (global batch B on one GPU) so that ncu can list / profile the launches:
    ncu --metrics gpu__time_duration.sum ... python tools/profile_train.py [B] [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import music_style_transfer_ldm_b200 as M
from music_style_transfer_ldm_b200.train import LDMTrainer
from oracle import ldm_oracle as O

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda")
m = M.LDM(32, pretrained_path=None)
m.load_state_dict(O.make_state_dict(1234))
m._freeze_encoder()
m.to(dev).set_precision("bf16")
tr = LDMTrainer(m, None, dev, lr=5e-4, style_loss_weight=3.0)
m.train()
g = torch.Generator().manual_seed(1)
x, s = torch.rand((B, 1, 128, 128), generator=g).to(dev), torch.rand((B, 1, 128, 128), generator=g).to(dev)
for i in range(steps):
    out = tr.train_step_device(x, s)
torch.cuda.synchronize()
print("ok", float(out["total_loss"]))
