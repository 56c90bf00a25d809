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


AE_LOSS_RTOL = (2e-5, 1e-4, 2e-3)      # the KL gradient's conditioning (make_golden_ae.py) shows from the second update on


def ae_grad_tol(key: str) -> float:
    """max-norm relative tolerance of a step-0 gradient: the decoder sees only MSE(recon, x); the encoder also the
    ill-conditioned KL term"""
    return 1e-4 if key.startswith("decoder.") else 1e-2


def test_autoencoder_steps_match_reference(sd):
    """oracle/train_oracle.py: OracleAutoencoderTrainer against the fixture made from the reference's encoder / decoder /
    compression_loss / AdamW (tests/golden/make_golden_ae.py)."""
    import make_golden_ae as GA
    g = cases.load("ae_steps")
    tr = T.OracleAutoencoderTrainer(sd, lr=GA.LR)
    for i in range(GA.STEPS):
        np.testing.assert_allclose(tr.train_step(GA.step_input(i)), g["losses"][i], rtol=AE_LOSS_RTOL[i], err_msg=f"step {i}")
        if i == 0:
            checked = 0
            for k in GA.SAVE_KEYS:
                if "g0:" + k in g:
                    want, got = g["g0:" + k], tr.sd[k].grad.numpy()
                    assert np.abs(got - want).max() <= ae_grad_tol(k) * np.abs(want).max(), k
                    checked += 1
            assert checked >= 8
    after = tr.state_dict()
    for k in GA.SAVE_KEYS:                  # three Adam steps move a weight by at most ~3 lr
        np.testing.assert_allclose(after[k].numpy(), g["p:" + k], rtol=1e-3, atol=7 * GA.LR, err_msg=k)
