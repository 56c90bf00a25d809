#!/usr/bin/env python
"""Development probe: isolated time of the attention core at the UNet's shapes (graph-replayed, CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_style_transfer_ldm_b200 import ops
for (b, l, s, e) in [(64, 16, 16, 256), (64, 4, 4, 512), (256, 16, 16, 256)]:
    q = torch.randn((b, l, e), device="cuda").bfloat16()
    kv = torch.randn((b, s, 2 * e), device="cuda").bfloat16()
    for _ in range(3): ops.attention(q, kv[:, :, :e], kv[:, :, e:], 4)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(50): o = ops.attention(q, kv[:, :, :e], kv[:, :, e:], 4)
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): g.replay()
    e1.record(); torch.cuda.synchronize()
    print(f"attention b{b} l{l} s{s} e{e}: {1e3 * e0.elapsed_time(e1) / 500:6.2f} us per launch")
