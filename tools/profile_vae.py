#!/usr/bin/env python
"""Config-2 VAE round trip (batch 16, 128 x 1024), eager, for ncu launch lists:
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/profile_vae.py [eval|train] [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import music_style_transfer_ldm_b200 as M
from oracle import ldm_oracle as O      # weights only (tooling, not product)

mode = sys.argv[1] if len(sys.argv) > 1 else "eval"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
sd = O.make_state_dict(1234)
enc, dec = M.SpectrogramEncoder(32), M.SpectrogramDecoder(32)
enc.load_state_dict({k[8:]: v for k, v in sd.items() if k.startswith("encoder.")})
dec.load_state_dict({k[8:]: v for k, v in sd.items() if k.startswith("decoder.")})
enc.cuda().set_precision("bf16"); dec.cuda().set_precision("bf16")
enc.train(mode == "train"); dec.train(mode == "train")
x = torch.rand((16, 1, 128, 1024), generator=torch.Generator().manual_seed(3)).cuda()
with torch.no_grad():
    for _ in range(reps):
        y = dec(enc(x))
torch.cuda.synchronize()
print("ok", tuple(y.shape), float(y.mean()))
