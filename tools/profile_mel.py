#!/usr/bin/env python
"""Runs the mel front end on 64 clips of 10 s a few times (ncu target: -k regex:mel_power_kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_style_transfer_ldm_b200.audio_processor import AudioPreprocessor

ap = AudioPreprocessor()
w = (0.1 * torch.randn((64, 220500), generator=torch.Generator().manual_seed(1))).cuda()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    x = ap.spectrogram_batch(w, n_mels=128, crop_frames=384)
torch.cuda.synchronize()
print("ok", float(x.mean()))
