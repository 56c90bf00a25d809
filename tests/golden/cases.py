"""Seeded inputs for the golden cases (shared by make_golden.py and the tests).

Inputs are regenerated from CPU ``torch.Generator`` seeds instead of being stored;
each fixture carries a float64 checksum of its inputs so drift in the generator
would be reported as such, not as a parity failure.
"""
from __future__ import annotations

import os

import numpy as np
import torch

GOLDEN_DIR = os.path.dirname(os.path.abspath(__file__))
WEIGHT_SEED = 1234


def _g(seed):
    return torch.Generator().manual_seed(seed)


def spec(seed, b, h=128, w=128):
    """Spectrogram-like input in [0,1] (dataset contract, dataset.py:51-53)."""
    return torch.rand((b, 1, h, w), generator=_g(seed))


def latent(seed, b, h=16, w=16, c=32):
    return torch.randn((b, c, h, w), generator=_g(seed))


def style_dict(seed, b, h=16, w=16):
    """Random post-ReLU-like style maps at the two resolutions the UNet consumes."""
    g = _g(seed)
    return {"s5": torch.rand((b, 256, h // 4, w // 4), generator=g),
            "s6": torch.rand((b, 512, h // 8, w // 8), generator=g)}


def checksum(*tensors) -> float:
    return float(sum(t.double().sum().item() + (t.double() ** 2).sum().item() for t in tensors))


def load(name):
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    with np.load(path) as z:
        return {k: z[k] for k in z.files}


CASES = {
    # name: dict of parameters consumed by make_golden.py and the tests
    "unet_c1": dict(b=1, t=[199]),                                   # BASELINE config 1
    "unet_b3": dict(b=3, t=[0, 57, 199]),
    "unet_wide": dict(b=1, t=[120], h=16, w=32),                      # 128x256 spectrogram latent
    "style_enc": dict(b=2),
    "vae_train": dict(b=2, training=True),
    "vae_eval": dict(b=2, training=False),
    "vae_wide_eval": dict(b=1, training=False, h=128, w=256),
    "ldm_forward": dict(b=2, t=[13, 180]),
    "transfer_b1": dict(b=1, n=8, eta=0.0),                           # unpatched reference
    "transfer_b2": dict(b=2, n=10, eta=0.5),                          # reference with t.item() neutralised
    "style_only_b1": dict(b=1, n=6, eta=0.0),
    "losses": dict(b=2),
}
