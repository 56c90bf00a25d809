#!/usr/bin/env python
"""Times the VGGish conv stack (768 images) with the current MSTK_CONV_HALO / MSTK_HALO_GS environment."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_style_transfer_ldm_b200 import ops

NIMG = int(sys.argv[1]) if len(sys.argv) > 1 else 768
LAYERS = [(NIMG, 64, 64, 64, 128), (NIMG, 32, 32, 128, 256), (NIMG, 32, 32, 256, 256), (NIMG, 16, 16, 256, 512), (NIMG, 16, 16, 512, 512)]
tot = 0.0
for n, h, w, cin, cout in LAYERS:
    x = torch.randn((n, h, w, cin), device="cuda").bfloat16()
    wp = torch.randn((9, cout, cin), device="cuda").bfloat16() * 0.05
    b = torch.zeros(cout, device="cuda")
    f = lambda: ops.conv(x, wp, b, ksize=3, act=ops.ACT_RELU)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): f()
    e1.record(); torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / 5
    fl = 2.0 * n * h * w * cin * cout * 9
    tot += us
    print(f"{us:9.1f} us {fl / us / 1e6:8.1f} TF  conv k3 {cin}->{cout} @{h}x{w} n{n}")
print(f"total {tot:.1f} us  [halo {os.environ.get('MSTK_CONV_HALO', '1')} gs {os.environ.get('MSTK_HALO_GS', '3,2')}]")
