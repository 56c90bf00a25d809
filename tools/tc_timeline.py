#!/usr/bin/env python
"""Development aid: per-phase clock64 timeline of CTA (0,0,0) of the tcgen05 conv kernel for the
UNet's layer shapes at batch 64 (stamps: 0 entry, 1 setup done, 2 first TMA issued, 3 first
stage landed, 4 last stage landed, 5 accumulator ready, 6 epilogue done, 7 dealloc; 8-10 split-K reduction;
11-15 epilogue phases; 16+8i start of k-block group i).  TCN=<batch> and TCVGG=1 select other shapes."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_style_transfer_ldm_b200 import ops, _lib
lib = _lib.load()
lib.mstk_debug_tc_timeline.argtypes = [ctypes.c_void_p]
dbg = torch.zeros(96, dtype=torch.int64, device="cuda")
lib.mstk_debug_tc_timeline(dbg.data_ptr())
CASES = [(64, 16, 16, 32, 64, 3, 1, False), (64, 16, 16, 64, 128, 3, 2, False), (64, 8, 8, 128, 256, 3, 2, False),
         (64, 4, 4, 256, 256, 1, 1, False), (64, 4, 4, 256, 512, 3, 2, False), (64, 2, 2, 512, 512, 3, 1, False),
         (64, 2, 2, 512, 256, 3, 2, True), (64, 8, 8, 128, 64, 3, 2, True), (64, 16, 16, 64, 32, 3, 1, False)]
if os.environ.get('TCVGG'):
    CASES = [(768, 64, 64, 64, 128, 3, 1, False), (768, 32, 32, 128, 256, 3, 1, False), (768, 16, 16, 512, 512, 3, 1, False)]
for (n, h, w, cin, cout, k, s, tr) in CASES:
    x = torch.randn((n, h, w, cin), device="cuda").bfloat16()
    wp = torch.randn((k * k, cout, cin), device="cuda").bfloat16()
    b = torch.zeros(cout, device="cuda")
    for _ in range(3):
        ops.conv(x, wp, b, ksize=k, stride=s, transposed=tr, act=1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.conv(x, wp, b, ksize=k, stride=s, transposed=tr, act=1); e1.record()
    torch.cuda.synchronize()
    t = dbg.cpu().tolist()
    rel = [v - t[0] for v in t]
    print(f"{'convT' if tr else 'conv'} k{k}s{s} {cin}->{cout} @{h}x{w}: event {1e3*e0.elapsed_time(e1):7.1f} us | cycles since entry: {rel[1:8]}")
    tops = [rel[16 + it * 8] for it in range(8) if t[16 + it * 8]]
    if t[8] and t[9]:
        print(f"    split-K: partial staged {rel[8]}, reduced into L2 + counted {rel[9]}, last CTA reloaded + scratch zeroed {rel[10]}")
    print(f"    epilogue: accumulator ready {rel[5]}, first tmem_ld {rel[13]}, first chunk math {rel[14]}, first chunk staged {rel[15]}, all staged {rel[11]}, store issued {rel[12]}, done {rel[6]}")
    print("    k-block start (after its wait), deltas:", [b - a for a, b in zip(tops, tops[1:])])
    dbg.zero_()
