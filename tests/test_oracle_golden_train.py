"""Pins oracle/train_oracle.py (losses of 3 consecutive steps and the updated parameters) against the fixture
produced by the unmodified reference LDMTrainer (tests/golden/make_golden_train.py).  CPU only."""
import numpy as np
import torch

import cases
import make_golden_train as G
from oracle import train_oracle as T


def test_train_steps_match_reference(sd):
    g = cases.load("train_steps")
    tr = T.OracleTrainer(sd, lr=G.LR, style_loss_weight=G.W)
    for i in range(G.STEPS):
        x, s, t, eps = G.step_inputs(i)
        out = tr.train_step(x, s, t, eps)
        got = [out["compression_loss"], out["denoisinsg_loss"], out["style_loss"], out["total_loss"]]
        np.testing.assert_allclose(got, g["losses"][i], rtol=2e-5, atol=1e-6, err_msg=f"step {i}")
    after = tr.state_dict()
    for k in G.SAVE_KEYS:
        # Adam's first steps move every weight by ~lr regardless of the gradient's size, so tiny summation-order
        # differences in a gradient that is ~0 can flip a step: compare with an absolute tolerance of a fraction of lr
        np.testing.assert_allclose(after[k].numpy(), g["p:" + k], rtol=1e-4, atol=2e-5, err_msg=k)
