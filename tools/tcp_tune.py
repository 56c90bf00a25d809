#!/usr/bin/env python
"""Times the throughput-bound tcgen05 layers (VGGish stack at 768 images, config-2 VAE layers) with the current
MSTK_TC_PERSIST / MSTK_TCP_FORCE environment: one line per layer (us, TFLOP/s)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_style_transfer_ldm_b200 import ops

LAYERS = [  # n, h, w, cin, cout, k, stride, transposed
    (768, 64, 64, 64, 128, 3, 1, False), (768, 32, 32, 128, 256, 3, 1, False), (768, 32, 32, 256, 256, 3, 1, False),
    (768, 16, 16, 256, 512, 3, 1, False), (768, 16, 16, 512, 512, 3, 1, False),
    (16, 64, 512, 64, 128, 3, 2, False), (16, 16, 128, 32, 128, 4, 2, True), (16, 32, 256, 128, 64, 4, 2, True)]
tot = 0.0
for n, h, w, cin, cout, k, s, tr in LAYERS:
    x = torch.randn((n, h, w, cin), device="cuda").bfloat16()
    wp = torch.randn((k * k, cout, cin), device="cuda").bfloat16() * 0.05
    b = torch.zeros(cout, device="cuda")
    f = lambda: ops.conv(x, wp, b, ksize=k, stride=s, transposed=tr, act=ops.ACT_RELU)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): f()
    e1.record(); torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / 5
    ho, wo = (2 * h, 2 * w) if tr else ((h + 2 - k) // s + 1 + (0 if k == 3 else 0), (w + 2 - k) // s + 1)
    fl = 2.0 * n * (h * w if tr else ho * wo) * cin * cout * k * k
    tot += us
    print(f"{us:9.1f} us {fl / us / 1e6:8.1f} TF  {'convT' if tr else 'conv'} k{k}s{s} {cin}->{cout} @{h}x{w} n{n}")
print(f"total {tot:.1f} us  [{os.environ.get('MSTK_TC_PERSIST', '1')} / {os.environ.get('MSTK_TCP_FORCE', '-')}]")
