"""100-step loss curve of the training step for the 1 % loss-curve parity test (north_star).

    python tests/golden/make_golden_curve.py   ->  tests/golden/train_curve_100.npz

Produced by oracle/train_oracle.py (fp32 CPU), which tests/test_oracle_golden_train.py pins to the unmodified
reference LDMTrainer on its first steps; running the reference itself for 100 steps gives the same numbers but the
reference tree is not available where the GPU tests run.  Batch 4, lr 5e-4, style_loss_weight 3.0 (config.py:3,9).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(os.path.dirname(HERE)), HERE]
import cases      # noqa: E402
from oracle import ldm_oracle as O, train_oracle as T   # noqa: E402

STEPS, B, LR, W = 100, 4, 5e-4, 3.0


def step_inputs(i):
    g = torch.Generator().manual_seed(9000 + i)
    return (torch.rand((B, 1, 128, 128), generator=g), torch.rand((B, 1, 128, 128), generator=g),
            torch.randint(0, 200, (B,), generator=g), torch.randn((B, 32, 16, 16), generator=g))


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    tr = T.OracleTrainer(O.make_state_dict(cases.WEIGHT_SEED), lr=LR, style_loss_weight=W)
    rows = []
    for i in range(STEPS):
        out = tr.train_step(*step_inputs(i))
        rows.append([out["compression_loss"], out["denoisinsg_loss"], out["style_loss"], out["total_loss"]])
        if i % 10 == 0:
            print(i, rows[-1], flush=True)
    np.savez_compressed(os.path.join(HERE, "train_curve_100.npz"), losses=np.asarray(rows, dtype=np.float64))


if __name__ == "__main__":
    main()
