#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference modules on CPU.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py

The reference holds no golden vectors of its own, so these
fixtures are "outputs of the reference itself run here": weights from
oracle.ldm_oracle.make_state_dict(WEIGHT_SEED) loaded with strict=True into the
reference LDM, inputs from tests/golden/cases.py, fp32, CPU.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import cases  # noqa: E402
import refshim  # noqa: E402
from oracle import ldm_oracle as O  # noqa: E402


def save(name, **arrays):
    out = {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in arrays.items()}
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"{name}: " + ", ".join(f"{k}{tuple(v.shape)}" for k, v in out.items()))


class fixed_randn_like:
    """Make the reference's internal ``torch.randn_like`` (model.py:110) return a
    pre-drawn tensor so the same noise can be fed to the oracle and the CUDA path."""

    def __init__(self, eps):
        self.eps = eps

    def __enter__(self):
        self._real = torch.randn_like
        torch.randn_like = lambda x, **kw: self.eps.to(x.dtype)
        return self

    def __exit__(self, *exc):
        torch.randn_like = self._real
        return False


class fixed_randn:
    def __init__(self, z):
        self.z = z

    def __enter__(self):
        self._real = torch.randn
        torch.randn = lambda *a, **kw: self.z.clone()
        return self

    def __exit__(self, *exc):
        torch.randn = self._real
        return False


@torch.no_grad()
def main():
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    sd = O.make_state_dict(cases.WEIGHT_SEED)
    ldm = refshim.build_ldm(sd)
    ldm.train()  # the reference's scripts never call .eval() on a freshly built LDM

    # ---- UNet eps-prediction (config 1 and friends) ----
    for name in ("unet_c1", "unet_b3", "unet_wide"):
        c = cases.CASES[name]
        h, w = c.get("h", 16), c.get("w", 16)
        z = cases.latent(11, c["b"], h, w)
        st = cases.style_dict(12, c["b"], h, w)
        t = torch.tensor(c["t"], dtype=torch.long)
        out = ldm.unet(z, t, st)
        save(name, eps_hat=out, in_checksum=cases.checksum(z, st["s5"], st["s6"]))

    # float timesteps, as models/tests.py:447 passes them
    z = cases.latent(11, 1)
    st = cases.style_dict(12, 1)
    out = ldm.unet(z, torch.tensor([37.0]), st)
    save("unet_float_t", eps_hat=out, in_checksum=cases.checksum(z, st["s5"], st["s6"]))

    # ---- style encoder ----
    c = cases.CASES["style_enc"]
    s = cases.spec(21, c["b"])
    emb = ldm.style_encoder(s)
    save("style_enc", s5=emb["s5"], s6=emb["s6"],
         tap_sums=np.array([emb[f"s{i}"].double().sum().item() for i in range(1, 7)]),
         tap_sqsums=np.array([(emb[f"s{i}"].double() ** 2).sum().item() for i in range(1, 7)]),
         in_checksum=cases.checksum(s))

    # ---- VAE encode / decode, train-mode and eval-mode BatchNorm ----
    for name in ("vae_train", "vae_eval", "vae_wide_eval"):
        c = cases.CASES[name]
        x = cases.spec(31, c["b"], c.get("h", 128), c.get("w", 128))
        ldm.encoder.train(c["training"])
        ldm.decoder.train(c["training"])
        # running buffers would be updated in train mode: restore afterwards
        z0 = ldm.encoder(x)
        y = ldm.decoder(z0)
        ldm.load_state_dict(sd, strict=True)
        save(name, z0=z0, decoded=y, in_checksum=cases.checksum(x))
    ldm.train()

    # ---- LDM.forward ----
    c = cases.CASES["ldm_forward"]
    x, s = cases.spec(41, c["b"]), cases.spec(42, c["b"])
    eps = cases.latent(43, c["b"])
    t = torch.tensor(c["t"], dtype=torch.long)
    with fixed_randn_like(eps):
        out = ldm(x, s, t)
    ldm.load_state_dict(sd, strict=True)
    assert torch.equal(out["noise"], eps)
    save("ldm_forward", z_t=out["z_t"], noise_pred=out["noise_pred"], z_0=out["z_0"],
         reconstructed=out["reconstructed"], in_checksum=cases.checksum(x, s, eps))
    fwd = out

    # ---- losses on the forward outputs (loss.py) ----
    L = refshim.load().loss
    den = L.diffusion_loss(fwd["noise_pred"], fwd["noise"])
    kl = L.kl_regularization_loss(fwd["z_0"])
    comp = L.compression_loss(x, fwd["reconstructed"], fwd["z_0"], ldm.feature_loss_net)
    sty = L.style_loss(fwd["reconstructed"], s, ldm.feature_loss_net)
    vgg = ldm.feature_loss_net(x, s)
    save("losses", diffusion=den, kl=kl, compression=comp, style=sty, vggish_x_s=vgg,
         total=comp + den + 3.0 * sty)

    # ---- content+style transfer (model.py:468-559) ----
    for name in ("transfer_b1", "transfer_b2"):
        c = cases.CASES[name]
        x, s = cases.spec(51, c["b"]), cases.spec(52, c["b"])
        eps = cases.latent(53, c["b"])
        with fixed_randn_like(eps), refshim.BatchedSamplingLDM():
            dec, ztd = ldm.content_style_transfer_wrapper(x, s, num_timesteps=c["n"], eta=c["eta"])
        ldm.load_state_dict(sd, strict=True)
        save(name, decoded=dec, z_t_decoded=ztd, in_checksum=cases.checksum(x, s, eps))

    # inner loop alone, with its per-step logs (B=1: unpatched reference)
    z_t = cases.latent(54, 1)
    st = cases.style_dict(55, 1)
    xs, logs = ldm.content_style_ddim_sample(z_t, st, timesteps=5, eta=0.25)
    save("ddim_inner", x=xs, timesteps=np.array(logs["timesteps"]),
         pred_x0=torch.stack(logs["pred_x0"]), noise_pred=torch.stack(logs["noise_pred"]),
         in_checksum=cases.checksum(z_t, st["s5"], st["s6"]))

    # ---- style-only generation (model.py:382-465) ----
    c = cases.CASES["style_only_b1"]
    s = cases.spec(61, c["b"])
    zT = cases.latent(62, c["b"])
    with fixed_randn(zT):
        dec = ldm.style_ddim_sample_wrapper((c["b"], 32, 16, 16), s, timesteps=c["n"], eta=c["eta"])
    ldm.load_state_dict(sd, strict=True)
    save("style_only_b1", decoded=dec, in_checksum=cases.checksum(s, zT))

    # ---- schedule buffers ----
    save("schedule", beta_t=ldm.noise_scheduler.beta_t, alpha_t=ldm.noise_scheduler.alpha_t,
         alpha_bar_t=ldm.noise_scheduler.alpha_bar_t,
         times_transfer_50=torch.linspace(49, 0, 50).long(),
         times_style_50=torch.linspace(199, 0, 50).long())


if __name__ == "__main__":
    main()
