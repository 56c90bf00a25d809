"""Functional fp32 CPU restatement of the reference's model / sampler / loss path.

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and
bench.py's baseline legs (the CPU arm, and -- tensors moved to `cuda` -- the
stock-PyTorch-on-this-GPU bar of SURVEY 8(d)).  Never by the product package.

Every function takes a flat ``sd`` (an ``LDM.state_dict()``-shaped mapping, keys
as in SURVEY.md Appendix A.1) so that the same weights can be pushed through
the reference, this oracle and the CUDA path.  The arithmetic substrate is
``torch.nn.functional`` on CPU, which is also what the reference runs on
(its ``nn.Module``s dispatch to the very same ATen ops), so pinning against the
reference is tight (<=1e-6 abs; see tests/test_oracle_golden.py).

Parity: PINNED by tests/golden/*.npz, produced by tests/golden/make_golden.py
from the unmodified reference modules.

Reference citations are to /root/reference/models/{model,loss}.py.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
SD = Dict[str, Tensor]

BN_EPS = 1e-5          # nn.BatchNorm2d default (model.py:18)
BN_MOMENTUM = 0.1


# --------------------------------------------------------------------------
# deterministic weights shared by reference / oracle / CUDA path
# --------------------------------------------------------------------------

def ldm_param_shapes(latent_dim: int = 32, num_timesteps: int = 200,
                     with_feature_net: bool = True) -> List[Tuple[str, Tuple[int, ...], str]]:
    """(key, shape, kind) for every entry of LDM.state_dict()."""
    out: List[Tuple[str, Tuple[int, ...], str]] = []

    def conv(name, co, ci, k):
        out.append((f"{name}.weight", (co, ci, k, k), "w"))
        out.append((f"{name}.bias", (co,), "b"))

    def convt(name, ci, co, k):
        out.append((f"{name}.weight", (ci, co, k, k), "wt"))
        out.append((f"{name}.bias", (co,), "b"))

    def bn(name, c):
        out.append((f"{name}.weight", (c,), "bn_w"))
        out.append((f"{name}.bias", (c,), "bn_b"))
        out.append((f"{name}.running_mean", (c,), "bn_rm"))
        out.append((f"{name}.running_var", (c,), "bn_rv"))
        out.append((f"{name}.num_batches_tracked", (), "bn_n"))

    conv("encoder.encoder.0", 64, 1, 3); bn("encoder.encoder.1", 64)
    conv("encoder.encoder.3", 128, 64, 3); bn("encoder.encoder.4", 128)
    conv("encoder.encoder.6", latent_dim, 128, 3); bn("encoder.encoder.7", latent_dim)
    convt("decoder.decoder.0", latent_dim, 128, 4); bn("decoder.decoder.1", 128)
    convt("decoder.decoder.3", 128, 64, 4); bn("decoder.decoder.4", 64)
    convt("decoder.decoder.6", 64, 1, 4)
    for i in (1, 3):
        out.append((f"unet.time_mlp.{i}.weight", (128, 128), "w"))
        out.append((f"unet.time_mlp.{i}.bias", (128,), "b"))
    conv("unet.enc1", 64, latent_dim, 3)
    conv("unet.enc2", 128, 64, 3)
    conv("unet.enc3", 256, 128, 3)
    conv("unet.enc4", 512, 256, 3)
    for name, e in (("unet.cross_attention1", 512), ("unet.cross_attention2", 256)):
        out.append((f"{name}.multihead_attn.in_proj_weight", (3 * e, e), "w"))
        out.append((f"{name}.multihead_attn.in_proj_bias", (3 * e,), "b"))
        out.append((f"{name}.multihead_attn.out_proj.weight", (e, e), "w"))
        out.append((f"{name}.multihead_attn.out_proj.bias", (e,), "b"))
    conv("unet.bottleneck", 512, 512, 3)
    convt("unet.dec4", 512, 256, 3)
    convt("unet.dec3", 256, 128, 3)
    convt("unet.dec2", 128, 64, 3)
    conv("unet.dec1", latent_dim, 64, 3)
    for k in ("beta_t", "alpha_t", "alpha_bar_t"):
        out.append((f"noise_scheduler.{k}", (num_timesteps,), "sched"))
    conv("style_encoder.enc1", 64, 1, 3)
    conv("style_encoder.enc2", 128, 64, 3)
    conv("style_encoder.enc3", 256, 128, 3)
    conv("style_encoder.enc4", 256, 256, 3)
    conv("style_encoder.enc5", 256, 256, 3)
    conv("style_encoder.enc6", 512, 256, 3)
    if with_feature_net:
        ci = 1
        for idx, co in ((0, 64), (3, 128), (6, 256), (8, 256), (11, 512), (13, 512)):
            conv(f"feature_loss_net.features.{idx}", co, ci, 3)
            ci = co
    return out


def noise_schedule(num_timesteps: int = 200) -> Tuple[Tensor, Tensor, Tensor]:
    """beta/alpha/alpha_bar exactly as ForwardDiffusion.__init__ (model.py:96-100)."""
    beta = torch.linspace(0.0001, 0.02, num_timesteps)
    alpha = 1 - beta
    return beta, alpha, torch.cumprod(alpha, dim=0)


def make_state_dict(seed: int = 0, latent_dim: int = 32, num_timesteps: int = 200,
                    with_feature_net: bool = True, bn_stats: bool = True) -> SD:
    """Seeded synthetic LDM weights (no checkpoint is obtainable offline).

    Scales follow PyTorch's default init magnitude (uniform +-1/sqrt(fan_in)) but
    the draw order is this function's own, so the dict is reproducible without the
    reference.  With ``bn_stats`` the BatchNorm affine/running buffers are made
    non-trivial so that eval-mode folding is actually exercised.
    """
    g = torch.Generator().manual_seed(seed)
    sd: SD = {}
    beta, alpha, abar = noise_schedule(num_timesteps)
    sched = {"beta_t": beta, "alpha_t": alpha, "alpha_bar_t": abar}
    for key, shape, kind in ldm_param_shapes(latent_dim, num_timesteps, with_feature_net):
        if kind in ("w", "wt"):
            if len(shape) == 4:
                fan_in = (shape[1] if kind == "w" else shape[0]) * shape[2] * shape[3]
                # ConvTranspose2d's default init uses weight.size(1)*k*k as fan_in
                if kind == "wt":
                    fan_in = shape[1] * shape[2] * shape[3]
            else:
                fan_in = shape[1]
            bound = 1.0 / math.sqrt(fan_in)
            sd[key] = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif kind == "b":
            sd[key] = (torch.rand(shape, generator=g) * 2 - 1) * 0.05
        elif kind == "bn_w":
            sd[key] = 1 + 0.2 * (torch.rand(shape, generator=g) - 0.5) if bn_stats else torch.ones(shape)
        elif kind == "bn_b":
            sd[key] = 0.1 * (torch.rand(shape, generator=g) - 0.5) if bn_stats else torch.zeros(shape)
        elif kind == "bn_rm":
            sd[key] = 0.1 * (torch.rand(shape, generator=g) - 0.5) if bn_stats else torch.zeros(shape)
        elif kind == "bn_rv":
            sd[key] = 0.5 + torch.rand(shape, generator=g) if bn_stats else torch.ones(shape)
        elif kind == "bn_n":
            sd[key] = torch.tensor(0, dtype=torch.long)
        elif kind == "sched":
            sd[key] = sched[key.split(".")[-1]].clone()
        else:  # pragma: no cover
            raise AssertionError(kind)
    return sd


# --------------------------------------------------------------------------
# building blocks
# --------------------------------------------------------------------------

def _bn(sd: SD, name: str, x: Tensor, training: bool, update: Optional[SD] = None) -> Tensor:
    """nn.BatchNorm2d forward; in training mode uses batch statistics (biased var)
    and, if ``update`` is given, writes the running-buffer update there."""
    w, b = sd[f"{name}.weight"], sd[f"{name}.bias"]
    if training:
        mean = x.mean(dim=(0, 2, 3))
        var = x.var(dim=(0, 2, 3), unbiased=False)
        if update is not None:
            n = x.numel() / x.shape[1]
            update[f"{name}.running_mean"] = (1 - BN_MOMENTUM) * sd[f"{name}.running_mean"] + BN_MOMENTUM * mean
            update[f"{name}.running_var"] = (1 - BN_MOMENTUM) * sd[f"{name}.running_var"] + BN_MOMENTUM * var * n / (n - 1)
            update[f"{name}.num_batches_tracked"] = sd[f"{name}.num_batches_tracked"] + 1
    else:
        mean, var = sd[f"{name}.running_mean"], sd[f"{name}.running_var"]
    inv = torch.rsqrt(var + BN_EPS)
    return (x - mean[None, :, None, None]) * (inv * w)[None, :, None, None] + b[None, :, None, None]


def encoder_fwd(sd: SD, x: Tensor, training: bool = False, prefix: str = "encoder.",
                update: Optional[SD] = None) -> Tensor:
    """SpectrogramEncoder.forward (model.py:16-28): 3x(conv3x3 s2) with BN, ReLU x2."""
    p = prefix + "encoder."
    h = F.conv2d(x, sd[p + "0.weight"], sd[p + "0.bias"], stride=2, padding=1)
    h = F.relu(_bn(sd, p + "1", h, training, update))
    h = F.conv2d(h, sd[p + "3.weight"], sd[p + "3.bias"], stride=2, padding=1)
    h = F.relu(_bn(sd, p + "4", h, training, update))
    h = F.conv2d(h, sd[p + "6.weight"], sd[p + "6.bias"], stride=2, padding=1)
    return _bn(sd, p + "7", h, training, update)


def decoder_fwd(sd: SD, z: Tensor, training: bool = False, prefix: str = "decoder.",
                update: Optional[SD] = None) -> Tensor:
    """SpectrogramDecoder.forward (model.py:37-49): 3x convT4x4 s2 p1, BN+ReLU x2, tanh."""
    p = prefix + "decoder."
    h = F.conv_transpose2d(z, sd[p + "0.weight"], sd[p + "0.bias"], stride=2, padding=1)
    h = F.relu(_bn(sd, p + "1", h, training, update))
    h = F.conv_transpose2d(h, sd[p + "3.weight"], sd[p + "3.bias"], stride=2, padding=1)
    h = F.relu(_bn(sd, p + "4", h, training, update))
    h = F.conv_transpose2d(h, sd[p + "6.weight"], sd[p + "6.bias"], stride=2, padding=1)
    return torch.tanh(h)


def style_encoder_fwd(sd: SD, s: Tensor, prefix: str = "style_encoder.") -> Dict[str, Tensor]:
    """StyleEncoder.forward (model.py:70-88): six conv3x3 s2 + ReLU, all taps returned."""
    out = {}
    h = s
    for i in range(1, 7):
        h = F.relu(F.conv2d(h, sd[f"{prefix}enc{i}.weight"], sd[f"{prefix}enc{i}.bias"], stride=2, padding=1))
        out[f"s{i}"] = h
    return out


def sinusoidal_embedding(t: Tensor, dim: int = 128) -> Tensor:
    """SinusoidalPositionEmbeddings.forward (model.py:239-246)."""
    half = dim // 2
    freq = torch.exp(torch.arange(half, device=t.device) * -(math.log(10000) / (half - 1)))
    arg = t[:, None] * freq[None, :]
    return torch.cat((arg.sin(), arg.cos()), dim=-1)


def time_mlp(sd: SD, t: Tensor, prefix: str = "unet.") -> Tensor:
    """UNet.time_mlp (model.py:170-175): sinusoid -> Linear -> GELU(erf) -> Linear."""
    e = sinusoidal_embedding(t, 128)
    e = F.linear(e, sd[prefix + "time_mlp.1.weight"], sd[prefix + "time_mlp.1.bias"])
    e = F.gelu(e)
    return F.linear(e, sd[prefix + "time_mlp.3.weight"], sd[prefix + "time_mlp.3.bias"])


def cross_attention(sd: SD, name: str, feat: Tensor, style: Tensor, num_heads: int = 4) -> Tensor:
    """CrossAttention.forward (model.py:135-160) with nn.MultiheadAttention math
    written out: q from UNet features, k/v from the style map, softmax
    over the style tokens of the same sample, out-proj; the result REPLACES feat."""
    B, C, H, W = feat.shape
    d = C // num_heads
    w_in, b_in = sd[name + ".multihead_attn.in_proj_weight"], sd[name + ".multihead_attn.in_proj_bias"]
    w_out, b_out = sd[name + ".multihead_attn.out_proj.weight"], sd[name + ".multihead_attn.out_proj.bias"]
    xq = feat.permute(0, 2, 3, 1).reshape(B, H * W, C)        # token order = row-major (h, w)
    xs = style.permute(0, 2, 3, 1).reshape(B, H * W, C)       # same H*W as feat (model.py:150)
    q = F.linear(xq, w_in[:C], b_in[:C]).view(B, -1, num_heads, d).transpose(1, 2)
    k = F.linear(xs, w_in[C:2 * C], b_in[C:2 * C]).view(B, -1, num_heads, d).transpose(1, 2)
    v = F.linear(xs, w_in[2 * C:], b_in[2 * C:]).view(B, -1, num_heads, d).transpose(1, 2)
    att = torch.softmax((q * (1.0 / math.sqrt(d))) @ k.transpose(-1, -2), dim=-1)
    o = (att @ v).transpose(1, 2).reshape(B, H * W, C)
    o = F.linear(o, w_out, b_out)
    return o.view(B, H, W, C).permute(0, 3, 1, 2)


def unet_fwd(sd: SD, z: Tensor, t: Tensor, style: Dict[str, Tensor], prefix: str = "unet.",
             taps: Optional[Dict[str, Tensor]] = None) -> Tensor:
    """UNet.forward (model.py:196-231).  Quirks kept: +t_emb AFTER ReLU(enc2); attention
    output replaces the features; dec4's skip is the PRE-attention z3; dec1 has no ReLU."""
    p = prefix
    temb = time_mlp(sd, t, p)[:, :, None, None]
    z1 = F.relu(F.conv2d(z, sd[p + "enc1.weight"], sd[p + "enc1.bias"], stride=1, padding=1))
    z2 = F.relu(F.conv2d(z1, sd[p + "enc2.weight"], sd[p + "enc2.bias"], stride=2, padding=1)) + temb
    z3 = F.relu(F.conv2d(z2, sd[p + "enc3.weight"], sd[p + "enc3.bias"], stride=2, padding=1))
    a3 = cross_attention(sd, p + "cross_attention2", z3, style["s5"])
    z4 = F.relu(F.conv2d(a3, sd[p + "enc4.weight"], sd[p + "enc4.bias"], stride=2, padding=1))
    a4 = cross_attention(sd, p + "cross_attention1", z4, style["s6"])
    bt = F.relu(F.conv2d(a4, sd[p + "bottleneck.weight"], sd[p + "bottleneck.bias"], stride=1, padding=1))
    d4 = F.relu(F.conv_transpose2d(bt, sd[p + "dec4.weight"], sd[p + "dec4.bias"], stride=2, padding=1,
                                   output_padding=1)) + z3
    d3 = F.relu(F.conv_transpose2d(d4, sd[p + "dec3.weight"], sd[p + "dec3.bias"], stride=2, padding=1,
                                   output_padding=1)) + z2
    d2 = F.relu(F.conv_transpose2d(d3, sd[p + "dec2.weight"], sd[p + "dec2.bias"], stride=2, padding=1,
                                   output_padding=1)) + z1
    out = F.conv2d(d2, sd[p + "dec1.weight"], sd[p + "dec1.bias"], stride=1, padding=1)
    if taps is not None:
        taps.update(z1=z1, z2=z2, z3=z3, a3=a3, z4=z4, a4=a4, bt=bt, d4=d4, d3=d3, d2=d2)
    return out


# --------------------------------------------------------------------------
# scheduler arithmetic and the two DDIM loops
# --------------------------------------------------------------------------

def q_sample(sd: SD, x0: Tensor, t: Tensor, eps: Tensor) -> Tensor:
    """ForwardDiffusion.forward (model.py:102-115) with the noise passed in
    (the reference draws it with randn_like; parity tests pre-draw it)."""
    ab = sd["noise_scheduler.alpha_bar_t"][t].view(-1, 1, 1, 1)
    return torch.sqrt(ab) * x0 + torch.sqrt(1 - ab) * eps


def predict_x0(sd: SD, z_t: Tensor, t: Tensor, eps_hat: Tensor) -> Tensor:
    """ForwardDiffusion.predict_start_from_noise (model.py:117-124)."""
    ab = sd["noise_scheduler.alpha_bar_t"][t].view(-1, 1, 1, 1)
    return (z_t - torch.sqrt(1 - ab) * eps_hat) / torch.sqrt(ab)


def ddim_update(sd: SD, x: Tensor, eps_hat: Tensor, t: int, t_next: int, eta: float) -> Tuple[Tensor, Tensor]:
    """One x_t -> x_{t_next} update, operation order as model.py:441-458 / 535-552.
    Note the reference's ``eta`` only rescales the predicted-noise term; no fresh noise
    is ever injected.  Returns (x_next, x0_pred)."""
    abar = sd["noise_scheduler.alpha_bar_t"]
    ab_t, ab_n = abar[t], abar[t_next]
    x0 = (x - torch.sqrt(1 - ab_t) * eps_hat) / torch.sqrt(ab_t)
    dir_t = torch.sqrt(1 - ab_t) * eps_hat
    dir_n = torch.sqrt(1 - ab_n) * eps_hat
    x_next = torch.sqrt(ab_n) * x0 + dir_n + eta * (dir_n - dir_t)
    return x_next, x0


def ddim_times(first: int, n: int) -> List[int]:
    """torch.linspace(first, 0, n).long() (model.py:420, 514)."""
    return torch.linspace(first, 0, n).long().tolist()


def ddim_sample(sd: SD, z_t: Tensor, style: Dict[str, Tensor], times: List[int], eta: float = 0.0,
                logs: Optional[dict] = None) -> Tensor:
    """Shared body of style_conditioned_ddim_sample / content_style_ddim_sample
    (model.py:409-465, 503-559): len(times)-1 UNet evaluations."""
    x = z_t
    B = z_t.shape[0]
    for i in range(len(times) - 1):
        t, t_next = times[i], times[i + 1]
        tb = torch.full((B,), t, dtype=torch.long, device=z_t.device)
        eps_hat = unet_fwd(sd, x, tb, style)
        x, x0 = ddim_update(sd, x, eps_hat, t, t_next, eta)
        if logs is not None:
            logs.setdefault("timesteps", []).append(t)
            logs.setdefault("pred_x0", []).append(x0.clone())
            logs.setdefault("noise_pred", []).append(eps_hat.clone())
    return x


def content_style_transfer(sd: SD, content: Tensor, style_spec: Tensor, eps: Tensor, num_timesteps: int = 250,
                           eta: float = 0.0, training: bool = True) -> Tuple[Tensor, Tensor]:
    """LDM.content_style_transfer_wrapper (model.py:468-501).  ``training`` is the module
    mode of encoder/decoder (BatchNorm statistics): the reference's scripts leave the
    freshly built model in train mode.  Returns (decoded in [0,1], z_t_decoded in [-1,1])."""
    z0 = encoder_fwd(sd, content.float(), training)
    t = torch.full((content.shape[0],), num_timesteps - 1, dtype=torch.long, device=content.device)
    z_t = q_sample(sd, z0, t, eps)
    style = style_encoder_fwd(sd, style_spec.float())
    x = ddim_sample(sd, z_t, style, ddim_times(num_timesteps - 1, num_timesteps), eta)
    decoded = (decoder_fwd(sd, x, training) + 1) / 2
    return decoded, decoder_fwd(sd, z_t, training)


def style_only_generation(sd: SD, z_T: Tensor, style_spec: Tensor, timesteps: int = 100, eta: float = 0.0,
                          training: bool = True, num_train_timesteps: int = 200) -> Tensor:
    """LDM.style_ddim_sample_wrapper (model.py:382-407) with z_T passed in."""
    style = style_encoder_fwd(sd, style_spec)
    x = ddim_sample(sd, z_T, style, ddim_times(num_train_timesteps - 1, timesteps), eta)
    return (decoder_fwd(sd, x, training) + 1) / 2


def ldm_forward(sd: SD, x: Tensor, style_spec: Tensor, t: Tensor, eps: Tensor, training: bool = True,
                enc_training: Optional[bool] = None) -> Dict[str, Tensor]:
    """LDM.forward (model.py:355-379) with the q-sample noise passed in."""
    enc_training = training if enc_training is None else enc_training
    z0 = encoder_fwd(sd, x.float(), enc_training)
    style = style_encoder_fwd(sd, style_spec.float())
    z_t = q_sample(sd, z0, t, eps)
    eps_hat = unet_fwd(sd, z_t, t, style)
    z0_pred = predict_x0(sd, z_t, t, eps_hat)
    recon = (decoder_fwd(sd, z0_pred, training) + 1) / 2
    return {"z_t": z_t, "noise": eps, "noise_pred": eps_hat, "z_0": z0, "reconstructed": recon}


# --------------------------------------------------------------------------
# losses (models/loss.py)
# --------------------------------------------------------------------------

def diffusion_loss(eps_hat: Tensor, eps: Tensor) -> Tensor:
    """loss.py:48-49."""
    return torch.mean((eps_hat - eps) ** 2)


def kl_regularization_loss(latent: Tensor) -> Tensor:
    """loss.py:31-32."""
    sq = latent * latent
    return torch.mean(0.5 * (sq - 1 - torch.log(sq + 1e-8)))


VGGISH_CFG = [64, "M", 128, "M", 256, 256, "M", 512, 512, "M"]   # thirdparty/torchvggish/vggish.py:111
VGGISH_CONV_IDX = [0, 3, 6, 8, 11, 13]


def vggish_taps(sd: SD, x: Tensor, prefix: str = "feature_loss_net.features.") -> List[Tensor]:
    """The six post-ReLU feature maps of the VGGish conv stack (loss.py:78-90)."""
    taps = []
    idx = 0
    for v in VGGISH_CFG:
        if v == "M":
            x = F.max_pool2d(x, 2, 2)
            idx += 1
        else:
            x = F.relu(F.conv2d(x, sd[f"{prefix}{idx}.weight"], sd[f"{prefix}{idx}.bias"], padding=1))
            taps.append(x)
            idx += 2
    return taps


def vggish_feature_loss(sd: SD, pred: Tensor, target: Tensor) -> Tensor:
    """VGGishFeatureLoss.forward (loss.py:64-101): per tap, divide by the per-sample
    UNBIASED std over (C,H,W) + 1e-8, MSE, mean over the 6 taps.  No gradient."""
    total = 0.0
    tp, tt = vggish_taps(sd, pred), vggish_taps(sd, target)
    for a, b in zip(tp, tt):
        a = a / (torch.std(a, dim=[1, 2, 3], keepdim=True) + 1e-8)
        b = b / (torch.std(b, dim=[1, 2, 3], keepdim=True) + 1e-8)
        total = total + torch.mean((a - b) ** 2)
    return total / len(tp)


def compression_loss(sd: SD, original: Tensor, recon: Tensor, latent: Tensor) -> Tensor:
    """loss.py:34-45 on the 'vggish' branch (LPIPS: unavailable, parity unpinned)."""
    mse = torch.mean((recon - original) ** 2)
    return mse + 0.1 * vggish_feature_loss(sd, original, recon) + 0.01 * kl_regularization_loss(latent)


def style_loss(sd: SD, recon: Tensor, style_spec: Tensor) -> Tensor:
    """loss.py:104-106."""
    return vggish_feature_loss(sd, recon, style_spec)


def train_losses(sd: SD, content: Tensor, style_spec: Tensor, t: Tensor, eps: Tensor,
                 style_loss_weight: float = 3.0) -> Dict[str, Tensor]:
    """The four loss values of LDMTrainer.train_step (train.py:174-186), fp32, train-mode BN."""
    out = ldm_forward(sd, content, style_spec, t, eps, training=True)
    den = diffusion_loss(out["noise_pred"], out["noise"])
    comp = compression_loss(sd, content, out["reconstructed"], out["z_0"])
    sty = style_loss(sd, out["reconstructed"], style_spec)
    return {"compression_loss": comp, "denoisinsg_loss": den, "style_loss": sty,
            "total_loss": comp + den + style_loss_weight * sty}
