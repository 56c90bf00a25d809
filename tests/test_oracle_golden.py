"""Pins oracle/ldm_oracle.py against the fixtures produced by the unmodified
reference (tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

import cases
from oracle import ldm_oracle as O

ATOL = 2e-6   # fp32, same ATen kernels on both sides; differences are summation-order only


def close(a, b, atol=ATOL, rtol=1e-5):
    a = a.detach().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    np.testing.assert_allclose(a, b, atol=atol, rtol=rtol)


@pytest.mark.parametrize("name", ["unet_c1", "unet_b3", "unet_wide"])
def test_unet(sd, name):
    c = cases.CASES[name]
    g = cases.load(name)
    h, w = c.get("h", 16), c.get("w", 16)
    z, st = cases.latent(11, c["b"], h, w), cases.style_dict(12, c["b"], h, w)
    assert cases.checksum(z, st["s5"], st["s6"]) == pytest.approx(float(g["in_checksum"]), rel=1e-12)
    close(O.unet_fwd(sd, z, torch.tensor(c["t"]), st), g["eps_hat"])


def test_unet_float_timestep(sd):
    g = cases.load("unet_float_t")
    z, st = cases.latent(11, 1), cases.style_dict(12, 1)
    close(O.unet_fwd(sd, z, torch.tensor([37.0]), st), g["eps_hat"])


def test_style_encoder(sd):
    g = cases.load("style_enc")
    emb = O.style_encoder_fwd(sd, cases.spec(21, 2))
    close(emb["s5"], g["s5"])
    close(emb["s6"], g["s6"])
    for i in range(6):
        assert emb[f"s{i+1}"].double().sum().item() == pytest.approx(g["tap_sums"][i], rel=1e-5)


@pytest.mark.parametrize("name", ["vae_train", "vae_eval", "vae_wide_eval"])
def test_vae(sd, name):
    c = cases.CASES[name]
    g = cases.load(name)
    x = cases.spec(31, c["b"], c.get("h", 128), c.get("w", 128))
    z0 = O.encoder_fwd(sd, x, c["training"])
    close(z0, g["z0"], atol=1e-5)
    close(O.decoder_fwd(sd, z0, c["training"]), g["decoded"], atol=1e-5)


def test_ldm_forward_and_losses(sd):
    c = cases.CASES["ldm_forward"]
    g = cases.load("ldm_forward")
    x, s, eps = cases.spec(41, c["b"]), cases.spec(42, c["b"]), cases.latent(43, c["b"])
    out = O.ldm_forward(sd, x, s, torch.tensor(c["t"]), eps, training=True)
    for k in ("z_t", "noise_pred", "z_0", "reconstructed"):
        close(out[k], g[k], atol=1e-5)
    gl = cases.load("losses")
    close(O.diffusion_loss(out["noise_pred"], eps), gl["diffusion"], rtol=1e-5)
    close(O.kl_regularization_loss(out["z_0"]), gl["kl"], rtol=1e-5)
    close(O.compression_loss(sd, x, out["reconstructed"], out["z_0"]), gl["compression"], rtol=1e-4)
    close(O.style_loss(sd, out["reconstructed"], s), gl["style"], rtol=1e-4)
    close(O.vggish_feature_loss(sd, x, s), gl["vggish_x_s"], rtol=1e-4)
    tl = O.train_losses(sd, x, s, torch.tensor(c["t"]), eps)
    close(tl["total_loss"], gl["total"], rtol=1e-4)


@pytest.mark.parametrize("name", ["transfer_b1", "transfer_b2"])
def test_transfer(sd, name):
    c = cases.CASES[name]
    g = cases.load(name)
    x, s, eps = cases.spec(51, c["b"]), cases.spec(52, c["b"]), cases.latent(53, c["b"])
    dec, ztd = O.content_style_transfer(sd, x, s, eps, num_timesteps=c["n"], eta=c["eta"], training=True)
    close(dec, g["decoded"], atol=2e-5)
    close(ztd, g["z_t_decoded"], atol=2e-5)


def test_ddim_inner_logs(sd):
    g = cases.load("ddim_inner")
    z_t, st = cases.latent(54, 1), cases.style_dict(55, 1)
    logs = {}
    x = O.ddim_sample(sd, z_t, st, O.ddim_times(4, 5), eta=0.25, logs=logs)
    close(x, g["x"], atol=1e-5)
    assert logs["timesteps"] == g["timesteps"].tolist() == [4, 3, 2, 1]   # N-1 UNet calls
    close(torch.stack(logs["pred_x0"]), g["pred_x0"], atol=1e-5)
    close(torch.stack(logs["noise_pred"]), g["noise_pred"], atol=1e-5)


def test_style_only(sd):
    c = cases.CASES["style_only_b1"]
    g = cases.load("style_only_b1")
    s, zT = cases.spec(61, c["b"]), cases.latent(62, c["b"])
    close(O.style_only_generation(sd, zT, s, timesteps=c["n"], eta=c["eta"]), g["decoded"], atol=2e-5)


def test_schedule(sd):
    g = cases.load("schedule")
    beta, alpha, abar = O.noise_schedule(200)
    np.testing.assert_array_equal(beta.numpy(), g["beta_t"])
    np.testing.assert_array_equal(abar.numpy(), g["alpha_bar_t"])
    assert O.ddim_times(49, 50) == g["times_transfer_50"].tolist()
    assert O.ddim_times(199, 50) == g["times_style_50"].tolist()
    assert abar[0].item() == pytest.approx(0.9999, abs=1e-6) and abar[199].item() == pytest.approx(0.13218, abs=1e-4)


def test_state_dict_keys_match_reference_layout(sd):
    # shapes a reference checkpoint would carry
    assert sd["decoder.decoder.0.weight"].shape == (32, 128, 4, 4)
    assert sd["unet.dec4.weight"].shape == (512, 256, 3, 3)
    assert sd["unet.cross_attention1.multihead_attn.in_proj_weight"].shape == (1536, 512)
    n_unet = sum(v.numel() for k, v in sd.items() if k.startswith("unet."))
    assert n_unet == 6_841_504
    n_enc = sum(v.numel() for k, v in sd.items() if k.startswith("encoder.") and "running" not in k and "tracked" not in k)
    assert n_enc == 111_840
