"""Loss curve of the training step at BASELINE config 4's GLOBAL batch (256) for the first steps.

    python tests/golden/make_golden_curve256.py   ->  tests/golden/train_curve_b256.npz

Same recipe as make_golden_curve.py (oracle/train_oracle.py, fp32 CPU, pinned to the unmodified reference
LDMTrainer by tests/test_oracle_golden_train.py) at batch 256, lr 5e-4, style_loss_weight 3.0 (config.py:3,9),
3 steps: the fp32 mode is held to 1e-4 and the bf16 mode to 1 % on these, and bench.py's data-parallel
self-check (`ddp_parity`) compares the sharded step against the same numbers.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(os.path.dirname(HERE)), HERE]
import cases      # noqa: E402

STEPS, B, LR, W = 3, 256, 5e-4, 3.0


def step_inputs(i):
    g = torch.Generator().manual_seed(25600 + i)
    return (torch.rand((B, 1, 128, 128), generator=g), torch.rand((B, 1, 128, 128), generator=g),
            torch.randint(0, 200, (B,), generator=g), torch.randn((B, 32, 16, 16), generator=g))


def main():
    from oracle import ldm_oracle as O, train_oracle as T
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    tr = T.OracleTrainer(O.make_state_dict(cases.WEIGHT_SEED), lr=LR, style_loss_weight=W)
    rows = []
    for i in range(STEPS):
        out = tr.train_step(*step_inputs(i))
        rows.append([out["compression_loss"], out["denoisinsg_loss"], out["style_loss"], out["total_loss"]])
        print(i, rows[-1], flush=True)
    np.savez_compressed(os.path.join(HERE, "train_curve_b256.npz"), losses=np.asarray(rows, dtype=np.float64))


if __name__ == "__main__":
    main()
