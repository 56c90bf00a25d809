#!/usr/bin/env python
"""Per-op timeline of the layer-program kernel for a few UNet+DDIM steps at batch 64 (development aid)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import music_style_transfer_ldm_b200 as M
from music_style_transfer_ldm_b200 import _lib
from oracle import ldm_oracle as O

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
dev = "cuda"
m = M.LDM(32, pretrained_path=None)
m.load_state_dict(O.make_state_dict(1234))
m.to(dev).eval().set_precision("bf16")
g = torch.Generator().manual_seed(1)
z = torch.randn((64, 32, 16, 16), generator=g).to(dev)
st = {"s5": torch.rand((64, 256, 4, 4), generator=g).to(dev), "s6": torch.rand((64, 512, 2, 2), generator=g).to(dev)}
lib = _lib.load()
lib.mstk_debug_program_timeline.argtypes = [C.c_void_p]
lib.mstk_debug_tc_timeline.argtypes = [C.c_void_p]
tile_dbg = torch.zeros(16 * (steps * 20 + 8), dtype=torch.int64, device=dev)
lib.mstk_debug_tc_timeline(tile_dbg.data_ptr())       # recorded ops keep a pointer to their 16 stamps
with torch.no_grad():
    m.content_style_ddim_sample(z, st, timesteps=steps + 1, eta=0.0)      # builds + runs the program once
    lib.mstk_debug_tc_timeline(None)
    torch.cuda.synchronize()
    nops = list(m._programs.values())[0]['prog'].nops
    per = nops // steps
    buf = torch.zeros(2 * nops * 3, dtype=torch.int64, device=dev)
    lib.mstk_debug_program_timeline(buf.data_ptr())
    m.content_style_ddim_sample(z, st, timesteps=steps + 1, eta=0.0)
    torch.cuda.synchronize()
    lib.mstk_debug_program_timeline(None)
t = buf.cpu().view(2, nops, 3)
t0 = int(t[0, 0, 0])
print("op  | cta0: start  work  barrier | last cta: work barrier   (ns)")
for op in range(nops - per, nops):
    a, b = t[0, op], t[1, op]
    print(f"{op % per:3d} | {int(a[0]) - t0:8d} {int(a[1] - a[0]):6d} {int(a[2] - a[1]):6d} | {int(b[1] - b[0]):6d} {int(b[2] - b[1]):6d}")
print("last step total ns:", int(t[0, nops - 1, 1] - t[0, nops - per, 0]), 'ops/step', per)

print("per-op clock64 stamps of CTA 0 (cycles since stamp 0): 1 setup, 2 first TMA, 3 first stage landed, 4 last stage landed, 5 accum ready,")
print("  13 first tmem ld, 14 math, 15 staged, 11 epi loop done, 12 before TMA store, 6 epilogue end; 8/9/10 split-K stage/reduce/reload")
td = tile_dbg.cpu().view(-1, 16)
for op in range(nops - per, nops):
    r = td[op].tolist()
    if r[0] == 0:
        print(f"{op % per:3d} | (not a conv)")
        continue
    rel = {i: r[i] - r[0] for i in (1, 2, 3, 4, 5, 8, 9, 10, 13, 14, 15, 11, 12, 6) if r[i] >= r[0]}
    print(f"{op % per:3d} | " + " ".join(f"{k}:{v}" for k, v in rel.items()))
