"""CPU restatement of the reference's training step.  TEST INFRASTRUCTURE ONLY.

Follows models/train.py:163-208 (LDMTrainer.train_step) with the forward / losses of
oracle/ldm_oracle.py, autograd for the gradients and torch.optim.Adam for the update
(train.py:156: Adam over the parameters with requires_grad, default betas/eps, no weight decay).
The encoder is frozen as in the reference's LDM stage (model.py:340-341) but runs with train-mode
BatchNorm (train.py:212); the VGGish terms carry no gradient (loss.py:78, torch.no_grad()).
fp32 throughout (the reference's autocast/GradScaler is a precision choice, not part of the math).
``t`` and the q-sample noise are passed in (the reference draws them with torch.randint / randn_like).

Parity: PINNED by tests/golden/train_steps.npz, produced by tests/golden/make_golden_train.py from the
unmodified reference LDMTrainer (autocast neutralised, randint / randn_like fed the pre-drawn values).
"""
from __future__ import annotations

from typing import Dict

import torch

from . import ldm_oracle as O

TRAINABLE_PREFIXES = ("decoder.", "unet.", "style_encoder.")


def is_trainable(key: str) -> bool:
    return key.startswith(TRAINABLE_PREFIXES) and not key.endswith(("running_mean", "running_var", "num_batches_tracked"))


class OracleTrainer:
    def __init__(self, sd: O.SD, lr: float = 1e-4, style_loss_weight: float = 0.1):
        self.sd = {k: v.clone() for k, v in sd.items()}
        self.keys = [k for k in self.sd if is_trainable(k)]
        for k in self.keys:
            self.sd[k].requires_grad_(True)
        self.opt = torch.optim.Adam([self.sd[k] for k in self.keys], lr=lr)
        self.w = style_loss_weight

    def train_step(self, content: torch.Tensor, style: torch.Tensor, t: torch.Tensor, eps: torch.Tensor) -> Dict[str, float]:
        sd = self.sd
        self.opt.zero_grad()
        upd: O.SD = {}
        z0 = O.encoder_fwd(sd, content.float(), True, update=upd)
        emb = O.style_encoder_fwd(sd, style.float())
        z_t = O.q_sample(sd, z0, t, eps)
        eps_hat = O.unet_fwd(sd, z_t, t, emb)
        z0p = O.predict_x0(sd, z_t, t, eps_hat)
        rec = (O.decoder_fwd(sd, z0p, True, update=upd) + 1) / 2
        den = O.diffusion_loss(eps_hat, eps)
        with torch.no_grad():                      # loss.py:78: the feature stack runs without gradient
            perc = O.vggish_feature_loss(sd, content, rec)
            sty = O.vggish_feature_loss(sd, rec, style)
        comp = torch.mean((rec - content) ** 2) + 0.1 * perc + 0.01 * O.kl_regularization_loss(z0)
        total = comp + den + self.w * sty
        total.backward()
        self.opt.step()
        with torch.no_grad():
            for k, v in upd.items():               # BatchNorm running buffers (train mode)
                sd[k] = v.detach().clone() if isinstance(v, torch.Tensor) else v
        return {"compression_loss": float(comp), "denoisinsg_loss": float(den), "style_loss": float(sty), "total_loss": float(total)}

    def state_dict(self) -> O.SD:
        return {k: v.detach().clone() for k, v in self.sd.items()}


AE_PREFIXES = ("encoder.", "decoder.")


class OracleAutoencoderTrainer:
    """models/train.py:47, 68-83 (the autoencoder stage's loop body): latent = encoder(x); recon = decoder(latent)
    (tanh range, not rescaled); loss = compression_loss(x, recon, latent, VGGish) (loss.py:34-45, 'vggish' branch);
    torch.optim.AdamW over encoder + decoder parameters (default betas / eps / weight_decay 0.01).  Train-mode BatchNorm.

    Parity: PINNED by tests/golden/ae_steps.npz (tests/golden/make_golden_ae.py drives the unmodified reference
    SpectrogramEncoder / SpectrogramDecoder / compression_loss / VGGishFeatureLoss with the reference's optimizer):
    losses of three steps and the step-0 gradients.  The KL term's -1/z gradient at BatchNorm outputs near zero makes the
    encoder gradient ill-conditioned in the reference itself (see make_golden_ae.py), hence the looser encoder tolerances."""

    def __init__(self, sd: O.SD, lr: float = 5e-4):
        self.sd = {k: v.clone() for k, v in sd.items()}
        self.keys = [k for k in self.sd if k.startswith(AE_PREFIXES)
                     and not k.endswith(("running_mean", "running_var", "num_batches_tracked"))]
        for k in self.keys:
            self.sd[k].requires_grad_(True)
        self.opt = torch.optim.AdamW([self.sd[k] for k in self.keys], lr=lr)

    def train_step(self, x: torch.Tensor) -> float:
        sd = self.sd
        self.opt.zero_grad()
        upd: O.SD = {}
        z = O.encoder_fwd(sd, x.float(), True, update=upd)
        rec = O.decoder_fwd(sd, z, True, update=upd)
        with torch.no_grad():                      # loss.py:78
            perc = O.vggish_feature_loss(sd, x, rec)
        loss = torch.mean((rec - x) ** 2) + 0.1 * perc + 0.01 * O.kl_regularization_loss(z)
        loss.backward()
        self.opt.step()
        with torch.no_grad():
            for k, v in upd.items():
                sd[k] = v.detach().clone() if isinstance(v, torch.Tensor) else v
        return float(loss.detach())

    def state_dict(self) -> O.SD:
        return {k: v.detach().clone() for k, v in self.sd.items()}
