#!/usr/bin/env python
"""per-parameter gradient error of the autograd path vs the oracle (development aid): FP32_TC on / off"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")]
import torch
import cases, make_golden_train as G
from oracle import ldm_oracle as O
import test_gpu_autograd as TA
from music_style_transfer_ldm_b200 import ops
from gpu_util import rel_err
sd = O.make_state_dict(cases.WEIGHT_SEED)
x, s, t, eps = G.step_inputs(0)
ref, _ = TA._oracle_grads(sd, x, s, t, eps, G.W)
for tc in (True, False):
    ops.FP32_TC = tc
    m = TA._model(sd, "fp32")
    with TA._fixed_randn_like(eps):
        _, losses = TA._reference_style_losses(m, x.cuda(), s.cuda(), t.cuda(), G.W)
    losses["total_loss"].backward()
    named = dict(m.named_parameters())
    print("FP32_TC", tc)
    for k, g in ref.items():
        e = rel_err(named[k].grad, g)
        if e > 1e-3:
            d = (named[k].grad.cpu() - g).abs()
            print(f"  {k}: {e:.2e} of max norm {g.abs().max():.2e}; bad elements (>1e-2 of max) {(d > 1e-2 * g.abs().max()).sum().item()} / {g.numel()}")
