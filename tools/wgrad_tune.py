#!/usr/bin/env python
"""Times the weight-gradient kernel on the training step's layer shapes (current MSTK_WG_KSMAX): python tools/wgrad_tune.py [batch]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_style_transfer_ldm_b200 import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
# name, h, w (input), cin, cout, k, stride, transposed
LAYERS = [("bottleneck 512->512 k3 @2x2", 2, 2, 512, 512, 3, 1, False), ("enc4 256->512 k3s2 @4x4", 4, 4, 256, 512, 3, 2, False),
          ("dec4 convT 512->256 @2x2", 2, 2, 512, 256, 3, 2, True), ("attn1 kv 512->1024 k1 @2x2", 2, 2, 512, 1024, 1, 1, False),
          ("enc3 128->256 k3s2 @8x8", 8, 8, 128, 256, 3, 2, False), ("dec1 64->32 k3 @16x16", 16, 16, 64, 32, 3, 1, False),
          ("enc1 32->64 k3 @16x16", 16, 16, 32, 64, 3, 1, False), ("style enc3 128->256 k3s2 @32x32", 32, 32, 128, 256, 3, 2, False),
          ("style enc2 64->128 k3s2 @64x64", 64, 64, 64, 128, 3, 2, False), ("decoder d3 convT 128->64 k4 @32x32", 32, 32, 128, 64, 4, 2, True)]
tot = 0.0
for name, h, w, cin, cout, k, s, tr in LAYERS:
    x = torch.randn((B, h, w, cin), device="cuda").bfloat16()
    ho, wo = ops.conv_out_hw(h, w, k, s, tr)
    dy = torch.randn((B, ho, wo, cout), device="cuda").bfloat16()
    dw = torch.zeros((cin, cout, k, k) if tr else (cout, cin, k, k), device="cuda")
    f = lambda: ops.wgrad(x, dy, dw, ksize=k, stride=s, transposed=tr)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): f()
    e1.record(); torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / 10
    tot += us
    print(f"{us:8.1f} us  {name}")
print(f"total {tot:.1f} us  [batch {B}, MSTK_WG_KSMAX={os.environ.get('MSTK_WG_KSMAX', '-')}]")
