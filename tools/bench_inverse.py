#!/usr/bin/env python
"""Times the inverse audio path (64 images 128 x 128 -> 32 Griffin-Lim iterations -> waveforms) with CUDA events:
    python tools/bench_inverse.py [clips] [frames] [reps]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_style_transfer_ldm_b200.audio_processor import AudioPreprocessor
from music_style_transfer_ldm_b200 import ops

b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 128
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
ap = AudioPreprocessor()
g = torch.Generator().manual_seed(0)
img = (torch.rand((b, 128, frames), generator=g) * 255).to(torch.uint8).cuda()
u = torch.rand((b, frames, 1025), generator=g).cuda()
mag = ap.mel_images_to_magnitude(img)
out = {}
for name, fn in (("mel_to_stft", lambda: ap.mel_images_to_magnitude(img)), ("griffinlim32", lambda: ops.griffinlim(mag, u)),
                 ("griffinlim0", lambda: ops.griffinlim(mag, u, n_iter=0))):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    out[name + "_us"] = e0.elapsed_time(e1) * 1e3 / reps
out["clips"] = b; out["frames"] = frames
out["us_per_iteration"] = (out["griffinlim32_us"] - out["griffinlim0_us"]) / 32
# state streamed per iteration: angles r/w + tprev r/w (complex64) + S + windowed frames w + r = 8*4 + 4 + 8*2 per bin-ish
bytes_it = b * frames * (1025 * (8 * 4 + 4) + 2048 * 4 * 2)
out["algorithmic_gbs"] = bytes_it / out["us_per_iteration"] / 1e3
print(json.dumps(out))
