"""Drop-in module surface of the reference's models/model.py, executed by libmstk.

Constructor signatures, forward signatures, return structures and ``state_dict`` keys
follow the reference (file:line cited per class).  ``torch.nn`` layers are used ONLY as
parameter containers (so that keys, shapes and default initialisation are the
reference's); no ``forward`` of a torch layer is ever called -- every op is a kernel of
the C-ABI library reached through ``ops``.  There is no CPU path: tensors must be CUDA.

Precision: ``'fp32'`` (CUDA-core fp32 kernels, 1e-4 parity) or ``'bf16'`` (tcgen05
implicit-GEMM kernels with fp32 accumulation, 2e-2 parity); see ``set_precision``.
Activations between kernels are NHWC.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from . import ops
from .config import config
from .ops import ACT_NONE, ACT_RELU, ACT_TANH, MstkError

_DEFAULT_PRECISION = "fp32"


def set_default_precision(p: str) -> None:
    global _DEFAULT_PRECISION
    if p not in ("fp32", "bf16"):
        raise ValueError("precision must be 'fp32' or 'bf16'")
    _DEFAULT_PRECISION = p


def _act_dtype(precision: str) -> torch.dtype:
    return torch.bfloat16 if precision == "bf16" else torch.float32


class _Kernelised(nn.Module):
    """Mixin: precision switch + cache of packed parameters keyed on tensor versions."""

    precision: Optional[str] = None

    def set_precision(self, p: Optional[str]):
        if p not in (None, "fp32", "bf16"):
            raise ValueError("precision must be 'fp32', 'bf16' or None (inherit default)")
        for m in self.modules():
            if isinstance(m, _Kernelised) or hasattr(m, "_taps"):      # _taps: the VGGishFeatureLoss stack (loss.py)
                m.precision = p
        return self

    def _precision(self) -> str:
        return self.precision or _DEFAULT_PRECISION

    def _cached(self, key, deps, build):
        cache = self.__dict__.setdefault("_mstk_cache", {})
        sig = tuple((d.data_ptr(), d._version, d.device) for d in deps)
        hit = cache.get(key)
        if hit is not None and hit[0] == sig:
            return hit[1]
        with torch.no_grad():
            val = build()
        cache[key] = (sig, val)
        return val

    def _packed(self, name: str, w: torch.Tensor, dtype: torch.dtype, transposed: bool = False) -> torch.Tensor:
        return self._cached((name, dtype, transposed), (w,), lambda: ops.pack_weight(w, dtype, transposed=transposed))

    def _f32(self, name: str, t: torch.Tensor) -> torch.Tensor:
        """fp32 contiguous view of a bias-like parameter."""
        t = t.detach()
        if t.dtype == torch.float32 and t.is_contiguous():
            return t
        return self._cached((name, "f32"), (t,), lambda: t.float().contiguous())


def _guard(who, module, y, *inputs):
    from .autograd import guard
    return guard(who, module, y, *inputs)


def _need_cuda(x: torch.Tensor, who: str) -> None:
    if not x.is_cuda:
        raise MstkError(f"{who}: input is on {x.device}; this implementation runs on sm_100a kernels only (no CPU fallback)")


def _first_layer_conv(mod: _Kernelised, name: str, conv: nn.Conv2d, x_nhwc_f32: torch.Tensor, out_dtype, act, stats=None):
    """Cin == 1 stride-2 conv: HBM-bound, CUDA-core kernel, fp32 input and weights in both
    precisions, output in the activation dtype."""
    wp = mod._packed(name, conv.weight, torch.float32)
    return ops.conv(x_nhwc_f32, wp, mod._f32(name + ".b", conv.bias), ksize=3, stride=2, act=act, stats=stats,
                    out_dtype=out_dtype)


class _BnFold:
    """Eval-mode BatchNorm folded into the preceding conv: packed weight scaled per output
    channel and a fused bias (reference semantics: nn.BatchNorm2d in eval mode,
    model.py:18,21,24,39,42)."""

    @staticmethod
    def get(mod: _Kernelised, name: str, conv, bn: nn.BatchNorm2d, dtype, transposed: bool):
        deps = (conv.weight, conv.bias, bn.weight, bn.bias, bn.running_mean, bn.running_var)

        def build():
            scale = (bn.weight.detach() * torch.rsqrt(bn.running_var + bn.eps)).float().contiguous()
            bias = ((conv.bias.detach() - bn.running_mean) * scale + bn.bias.detach()).float().contiguous()
            wp = ops.pack_weight(conv.weight, dtype, transposed=transposed, scale=scale)
            return wp, bias

        return mod._cached((name, "fold", dtype), deps, build)


def _drop_fold(mod: _Kernelised, name: str) -> None:
    """a train-mode pass moves running_mean / running_var through raw pointers: the eval fold is stale"""
    cache = mod.__dict__.get("_mstk_cache")
    if cache:
        for k in [k for k in cache if isinstance(k, tuple) and len(k) == 3 and k[0] == name and k[1] == "fold"]:
            del cache[k]


def _conv_bn(mod: _Kernelised, name: str, conv, bn: nn.BatchNorm2d, x: torch.Tensor, *, w_dtype, mid_dtype, out_dtype,
             act, transposed: bool, ksize: int, first_layer: bool = False) -> torch.Tensor:
    """conv (or convT) + BatchNorm2d (+ReLU).  Train mode: batch statistics gathered in the
    conv epilogue, then one streaming normalise pass that also updates the running buffers.
    Eval mode: folded, single kernel."""
    if not bn.training:
        wp, bias = _BnFold.get(mod, name, conv, bn, w_dtype, transposed)
        return ops.conv(x, wp, bias, ksize=ksize, stride=2, transposed=transposed, act=act, out_dtype=out_dtype)
    cout = bn.num_features
    _drop_fold(mod, name)
    stats = ops.zeros_f64(2 * cout, x.device)
    wp = mod._packed(name, conv.weight, w_dtype, transposed)
    y = ops.conv(x, wp, mod._f32(name + ".b", conv.bias), ksize=ksize, stride=2, transposed=transposed, act=ACT_NONE,
                 stats=stats, out_dtype=mid_dtype)
    stats = _sync_bn_stats(stats)
    count = y.shape[0] * y.shape[1] * y.shape[2] * _bn_world()
    momentum = bn.momentum if bn.momentum is not None else 0.1
    track = bn.track_running_stats and bn.running_mean is not None
    out = ops.bn_apply(y, stats, mod._f32(name + ".g", bn.weight), mod._f32(name + ".be", bn.bias), count=count,
                       eps=bn.eps, momentum=momentum, act=act, out_dtype=out_dtype,
                       running_mean=bn.running_mean if track else None, running_var=bn.running_var if track else None)
    if track and bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)
    return out


# data-parallel hook: inside ``with bn_sync_group(group, world)`` the per-channel (sum, sumsq) of every
# train-mode BatchNorm are all-reduced so that each rank normalises with the GLOBAL batch statistics
# (SURVEY 8e).  The trainers open the scope around their own step only: a module forward outside a
# training step (a rank-0-only sample or validation pass) never issues a collective.
_BN_GROUP = {"group": None, "world": 1, "reduce": None}


class bn_sync_group:
    """``reduce``: optional callable that sums a [2C] fp64 statistics vector over the ranks in place (the trainer passes its
    peer-memory exchange); default: torch.distributed.all_reduce on ``group``."""

    def __init__(self, group, world_size: int, reduce=None):
        self.new = (group, int(world_size), reduce)

    def __enter__(self):
        self.old = (_BN_GROUP["group"], _BN_GROUP["world"], _BN_GROUP["reduce"])
        _BN_GROUP["group"], _BN_GROUP["world"], _BN_GROUP["reduce"] = self.new
        return self

    def __exit__(self, *exc):
        _BN_GROUP["group"], _BN_GROUP["world"], _BN_GROUP["reduce"] = self.old
        return False


def invalidate_packed(root: nn.Module, keep=()) -> None:
    """Parameters / buffers under ``root`` were rewritten through raw pointers (the trainers' fused Adam, a
    train-mode BatchNorm's running buffers), which torch's version counters do not see: drop every packed
    copy and every captured sampling graph that holds one."""
    for m in root.modules():
        if any(m is k for k in keep):
            continue
        m.__dict__.pop("_mstk_cache", None)
        m.__dict__.pop("_graphs", None)


def _bn_world() -> int:
    return _BN_GROUP["world"]


def _sync_bn_stats(stats: torch.Tensor) -> torch.Tensor:
    if _BN_GROUP["world"] > 1:
        if _BN_GROUP["reduce"] is not None:
            _BN_GROUP["reduce"](stats)
        else:
            import torch.distributed as dist
            dist.all_reduce(stats, group=_BN_GROUP["group"])
    return stats


class SpectrogramEncoder(_Kernelised):
    """reference models/model.py:10-28.  [B,1,H,W] -> [B,latent_dim,H/8,W/8]."""

    def __init__(self, latent_dim=4):
        super().__init__()
        self.encoder = nn.Sequential(
            nn.Conv2d(1, 64, kernel_size=3, stride=2, padding=1), nn.BatchNorm2d(64), nn.ReLU(),
            nn.Conv2d(64, 128, kernel_size=3, stride=2, padding=1), nn.BatchNorm2d(128), nn.ReLU(),
            nn.Conv2d(128, latent_dim, kernel_size=3, stride=2, padding=1), nn.BatchNorm2d(latent_dim))

    def _forward_nhwc(self, x: torch.Tensor) -> torch.Tensor:
        """x: [B,H,W,1] fp32 -> z0 [B,H/8,W/8,latent] fp32"""
        e = self.encoder
        adt = _act_dtype(self._precision())
        if e[1].training:
            cout = e[1].num_features
            _drop_fold(self, "c0")
            stats = ops.zeros_f64(2 * cout, x.device)
            y = _first_layer_conv(self, "c0", e[0], x, adt, ACT_NONE, stats=stats)
            stats = _sync_bn_stats(stats)
            bn = e[1]
            h = ops.bn_apply(y, stats, self._f32("g0", bn.weight), self._f32("b0", bn.bias),
                             count=y.shape[0] * y.shape[1] * y.shape[2] * _bn_world(), eps=bn.eps,
                             momentum=bn.momentum if bn.momentum is not None else 0.1, act=ACT_RELU, out_dtype=adt,
                             running_mean=bn.running_mean, running_var=bn.running_var)
            bn.num_batches_tracked.add_(1)
        else:
            wp, bias = _BnFold.get(self, "c0", e[0], e[1], torch.float32, False)
            h = ops.conv(x, wp, bias, ksize=3, stride=2, act=ACT_RELU, out_dtype=adt)
        h = _conv_bn(self, "c3", e[3], e[4], h, w_dtype=adt, mid_dtype=adt, out_dtype=adt, act=ACT_RELU,
                     transposed=False, ksize=3)
        return _conv_bn(self, "c6", e[6], e[7], h, w_dtype=adt, mid_dtype=torch.float32, out_dtype=torch.float32,
                        act=ACT_NONE, transposed=False, ksize=3)

    def forward(self, x):
        _need_cuda(x, "SpectrogramEncoder")
        x = x.detach().float().contiguous()
        if x.shape[1] != 1:
            raise MstkError("SpectrogramEncoder expects [B,1,H,W]")
        z = self._forward_nhwc(x.view(x.shape[0], x.shape[2], x.shape[3], 1))
        return _guard("SpectrogramEncoder", self, ops.nhwc_to_nchw(z), x)


class SpectrogramDecoder(_Kernelised):
    """reference models/model.py:31-49.  [B,latent_dim,h,w] -> [B,1,8h,8w] in [-1,1]."""

    def __init__(self, latent_dim=4):
        super().__init__()
        self.decoder = nn.Sequential(
            nn.ConvTranspose2d(latent_dim, 128, kernel_size=4, stride=2, padding=1), nn.BatchNorm2d(128), nn.ReLU(),
            nn.ConvTranspose2d(128, 64, kernel_size=4, stride=2, padding=1), nn.BatchNorm2d(64), nn.ReLU(),
            nn.ConvTranspose2d(64, 1, kernel_size=4, stride=2, padding=1), nn.Tanh())

    def _forward_nhwc(self, z: torch.Tensor, unit_range: bool = False) -> torch.Tensor:
        """z: [B,h,w,latent] in the activation dtype -> [B,8h,8w,1] fp32; ``unit_range`` fuses
        the callers' (y+1)/2 (model.py:371, 405, 498) into the tanh epilogue."""
        d = self.decoder
        adt = _act_dtype(self._precision())
        if z.dtype != adt:
            z = ops.cast(z, adt)
        h = _conv_bn(self, "d0", d[0], d[1], z, w_dtype=adt, mid_dtype=adt, out_dtype=adt, act=ACT_RELU,
                     transposed=True, ksize=4)
        h = _conv_bn(self, "d3", d[3], d[4], h, w_dtype=adt, mid_dtype=adt, out_dtype=adt, act=ACT_RELU,
                     transposed=True, ksize=4)
        wp = self._packed("d6", d[6].weight, adt, True)
        return ops.conv(h, wp, self._f32("d6.b", d[6].bias), ksize=4, stride=2, transposed=True, act=ACT_TANH,
                        out_dtype=torch.float32, out_scale=0.5 if unit_range else 1.0,
                        out_shift=0.5 if unit_range else 0.0)

    def forward(self, z):
        _need_cuda(z, "SpectrogramDecoder")
        zl = ops.nchw_to_nhwc(z.detach().float().contiguous(), _act_dtype(self._precision()))
        y = self._forward_nhwc(zl)
        return _guard("SpectrogramDecoder", self, y.view(y.shape[0], 1, y.shape[1], y.shape[2]), z)


class StyleEncoder(_Kernelised):
    """reference models/model.py:51-88.  Six conv3x3 s2 + ReLU; returns dict s1..s6 (NCHW)."""

    def __init__(self, in_channels=1, num_filters=64):
        super().__init__()
        nf = num_filters
        self.enc1 = nn.Conv2d(in_channels, nf, kernel_size=3, stride=2, padding=1)
        self.enc2 = nn.Conv2d(nf, nf * 2, kernel_size=3, stride=2, padding=1)
        self.enc3 = nn.Conv2d(nf * 2, nf * 4, kernel_size=3, stride=2, padding=1)
        self.enc4 = nn.Conv2d(nf * 4, nf * 4, kernel_size=3, stride=2, padding=1)
        self.enc5 = nn.Conv2d(nf * 4, nf * 4, kernel_size=3, stride=2, padding=1)
        self.enc6 = nn.Conv2d(nf * 4, nf * 8, kernel_size=3, stride=2, padding=1)

    def _forward_nhwc(self, s: torch.Tensor) -> List[torch.Tensor]:
        """s: [B,H,W,Cin] fp32 -> list of the six NHWC feature maps in the activation dtype"""
        adt = _act_dtype(self._precision())
        outs = []
        if self.enc1.in_channels == 1:
            h = _first_layer_conv(self, "enc1", self.enc1, s, adt, ACT_RELU)
        else:
            h = ops.conv(ops.cast(s, adt), self._packed("enc1", self.enc1.weight, adt), self._f32("enc1.b", self.enc1.bias),
                         ksize=3, stride=2, act=ACT_RELU)
        outs.append(h)
        for i in range(2, 7):
            c = getattr(self, f"enc{i}")
            h = ops.conv(h, self._packed(f"enc{i}", c.weight, adt), self._f32(f"enc{i}.b", c.bias), ksize=3, stride=2,
                         act=ACT_RELU)
            outs.append(h)
        return outs

    def forward(self, style_spectrogram):
        _need_cuda(style_spectrogram, "StyleEncoder")
        s = style_spectrogram.detach().float().contiguous()
        s_nhwc = s.view(s.shape[0], s.shape[2], s.shape[3], 1) if s.shape[1] == 1 else ops.nchw_to_nhwc(s, torch.float32)
        outs = self._forward_nhwc(s_nhwc)
        res = _StyleDict({f"s{i + 1}": ops.nhwc_to_nchw(o) for i, o in enumerate(outs)})
        res._nhwc = {f"s{i + 1}": o for i, o in enumerate(outs)}
        return _guard("StyleEncoder", self, res, style_spectrogram)


class _StyleDict(dict):
    """A plain dict (reference contract: keys s1..s6, NCHW fp32) that also carries the NHWC
    activations so that the UNet need not convert them back."""
    _nhwc: Dict[str, torch.Tensor] = {}


class ForwardDiffusion(_Kernelised):
    """reference models/model.py:90-124.  Linear beta schedule; buffers beta_t, alpha_t,
    alpha_bar_t live in the state_dict exactly as in the reference."""

    def __init__(self, num_timesteps=config["forward_diffusion_num_timesteps"]):
        super().__init__()
        self.num_timesteps = num_timesteps
        self.register_buffer("beta_t", torch.linspace(0.0001, 0.02, num_timesteps))
        self.register_buffer("alpha_t", 1 - self.beta_t)
        self.register_buffer("alpha_bar_t", torch.cumprod(self.alpha_t, dim=0))

    def _t(self, t: torch.Tensor, device) -> torch.Tensor:
        return t.to(device=device, dtype=torch.int64).contiguous()

    def forward(self, x_0, t):
        """(z_t, eps) with eps ~ N(0, I) drawn by torch's generator, as the reference does
        (model.py:110); only the arithmetic is a kernel."""
        _need_cuda(x_0, "ForwardDiffusion")
        x_0 = x_0.detach().float().contiguous()
        eps = torch.randn_like(x_0)
        z_t, _ = ops.q_sample(x_0, eps, self.alpha_bar_t, self._t(t, x_0.device))
        return z_t, eps

    def predict_start_from_noise(self, z_t, t, noise_pred):
        _need_cuda(z_t, "ForwardDiffusion")
        x0, _ = ops.predict_x0(z_t.detach().float().contiguous(), noise_pred.detach().float().contiguous(),
                               self.alpha_bar_t, self._t(t, z_t.device))
        return x0


class CrossAttention(_Kernelised):
    """reference models/model.py:126-160: nn.MultiheadAttention(embed_dim, num_heads) with
    Q from the UNet features and K = V from the style map; the output REPLACES the features
    (no residual, no norm).  Token order is row-major (h, w)."""

    def __init__(self, embed_dim, num_heads=4):
        super().__init__()
        self.multihead_attn = nn.MultiheadAttention(embed_dim, num_heads)
        self.embed_dim = embed_dim
        self.num_heads = num_heads

    # Two executions of the same layer (csrc/xattn.cu has the algebra):
    #  * composed (inference: module forward, UNet.forward, both sampling loops; E in {256, 512}): the style map goes
    #    through ONE k1 conv whose weight composes Wq^T Wk, Wo Wv and bq^T Wk per head -- once per sampling loop -- and
    #    every evaluation of the layer is a single `xattn` launch;
    #  * plain (the trainers, which need q / k / v for the parameter gradients; other E): K|V projection, q projection,
    #    attention core, out projection.
    def _composed_ok(self) -> bool:
        return self.embed_dim in (256, 512) and self.num_heads <= 16 and self.embed_dim % self.num_heads == 0

    def _project_style(self, style_nhwc: torch.Tensor) -> torch.Tensor:
        """The style side of the layer: [B,h,w,E] -> [B,h,w,ctot] (composed) or [B,h,w,2E] (K|V).  Depends only on
        the style, so the sampling loop hoists it out of the 49 steps."""
        e = self.embed_dim
        m = self.multihead_attn
        adt = style_nhwc.dtype
        if self._composed_ok():
            wp, bias = self._cached(("wcomp", adt), (m.in_proj_weight, m.in_proj_bias, m.out_proj.weight),
                                    lambda: ops.xattn_compose(m.in_proj_weight, m.in_proj_bias, m.out_proj.weight, self.num_heads, adt))
            return ops.conv(style_nhwc, wp, bias, ksize=1)
        return self._project_style_kv(style_nhwc)

    def _project_style_kv(self, style_nhwc: torch.Tensor) -> torch.Tensor:
        e = self.embed_dim
        m = self.multihead_attn
        adt = style_nhwc.dtype
        wkv = self._cached(("wkv", adt), (m.in_proj_weight,), lambda: ops.pack_weight(m.in_proj_weight.detach()[e:], adt))
        bkv = self._cached(("bkv",), (m.in_proj_bias,), lambda: m.in_proj_bias.detach()[e:].float().contiguous())
        return ops.conv(style_nhwc, wkv, bkv, ksize=1)

    def _forward_nhwc(self, feat: torch.Tensor, kv: torch.Tensor, kv_static: bool = False) -> torch.Tensor:
        e = self.embed_dim
        m = self.multihead_attn
        adt = feat.dtype
        b, h, w, _ = feat.shape
        s_tok = kv.shape[1] * kv.shape[2]
        if kv.shape[3] != 2 * e:          # composed projection
            o = ops.xattn(feat.view(b, h * w, e), kv.view(b, s_tok, kv.shape[3]), self._f32("bo", m.out_proj.bias), self.num_heads,
                          kv_static=kv_static)
            return o.view(b, h, w, e)
        wq = self._cached(("wq", adt), (m.in_proj_weight,), lambda: ops.pack_weight(m.in_proj_weight.detach()[:e], adt))
        bq = self._cached(("bq",), (m.in_proj_bias,), lambda: m.in_proj_bias.detach()[:e].float().contiguous())
        q = ops.conv(feat, wq, bq, ksize=1)
        kv2 = kv.view(b, s_tok, 2 * e)
        o = ops.attention(q.view(b, h * w, e), kv2[:, :, :e], kv2[:, :, e:], self.num_heads)
        wo = self._packed("wo", m.out_proj.weight, adt)
        return ops.conv(o.view(b, h, w, e), wo, self._f32("bo", m.out_proj.bias), ksize=1)

    def forward(self, unet_features, style_embedding):
        _need_cuda(unet_features, "CrossAttention")
        adt = _act_dtype(self._precision())
        if unet_features.shape[:2] != style_embedding.shape[:2] or unet_features[0, 0].numel() != style_embedding[0, 0].numel():
            # the reference reshapes the style map with the feature map's h*w (model.py:149-150): any map with that many
            # tokens passes there (tokens in row-major order), anything else is a reshape error
            raise MstkError("CrossAttention: the style map must hold as many tokens (h*w) as the feature map, same batch and channels")
        f = ops.nchw_to_nhwc(unet_features.detach().float().contiguous(), adt)
        s = ops.nchw_to_nhwc(style_embedding.detach().float().contiguous(), adt)
        return _guard("CrossAttention", self, ops.nhwc_to_nchw(self._forward_nhwc(f, self._project_style(s))), unet_features, style_embedding)


class SinusoidalPositionEmbeddings(_Kernelised):
    """reference models/model.py:234-246.  Stand-alone use only; inside the UNet the sinusoid
    is fused with the two Linear layers in one kernel."""

    def __init__(self, dim):
        super().__init__()
        self.dim = dim

    def forward(self, time):
        # a [dim] table per sample: tiny, computed once per call on the host device
        # (kept for API completeness; the hot path uses ops.time_embed)
        half = self.dim // 2
        coef = math.log(10000) / (half - 1)
        freq = torch.exp(torch.arange(half, device=time.device) * -coef)
        arg = time[:, None] * freq[None, :]
        return torch.cat((arg.sin(), arg.cos()), dim=-1)


class UNet(_Kernelised):
    """reference models/model.py:163-231.  eps-prediction network: conv / conv-transpose
    stages with fused bias+ReLU(+t_emb)(+skip) epilogues and two style cross-attention
    layers.  Quirks kept: +t_emb AFTER ReLU(enc2); the attention output replaces
    the features; dec4's skip is the PRE-attention z3; dec1 has no activation."""

    def __init__(self, in_channels=1, out_channels=1, num_filters=64):
        super().__init__()
        time_emb_dim = 128
        nf = num_filters
        self.time_mlp = nn.Sequential(SinusoidalPositionEmbeddings(time_emb_dim), nn.Linear(time_emb_dim, time_emb_dim),
                                      nn.GELU(), nn.Linear(time_emb_dim, time_emb_dim))
        self.enc1 = nn.Conv2d(in_channels, nf, kernel_size=3, stride=1, padding=1)
        self.enc2 = nn.Conv2d(nf, nf * 2, kernel_size=3, stride=2, padding=1)
        self.enc3 = nn.Conv2d(nf * 2, nf * 4, kernel_size=3, stride=2, padding=1)
        self.enc4 = nn.Conv2d(nf * 4, nf * 8, kernel_size=3, stride=2, padding=1)
        self.cross_attention1 = CrossAttention(embed_dim=512, num_heads=4)
        self.cross_attention2 = CrossAttention(embed_dim=256, num_heads=4)
        self.bottleneck = nn.Conv2d(nf * 8, nf * 8, kernel_size=3, stride=1, padding=1)
        self.dec4 = nn.ConvTranspose2d(nf * 8, nf * 4, kernel_size=3, stride=2, padding=1, output_padding=1)
        self.dec3 = nn.ConvTranspose2d(nf * 4, nf * 2, kernel_size=3, stride=2, padding=1, output_padding=1)
        self.dec2 = nn.ConvTranspose2d(nf * 2, nf, kernel_size=3, stride=2, padding=1, output_padding=1)
        self.dec1 = nn.Conv2d(nf, out_channels, kernel_size=3, stride=1, padding=1)

    # -- pieces used by both the module forward and the sampling loop ------------------
    def _time_embedding(self, t: torch.Tensor) -> torch.Tensor:
        """t: [B] int64 or float (models/tests.py:447 passes floats) -> [B,128] fp32"""
        l1, l2 = self.time_mlp[1], self.time_mlp[3]
        return ops.time_embed(t.detach().to(torch.float32).contiguous(), self._f32("t1w", l1.weight), self._f32("t1b", l1.bias),
                              self._f32("t2w", l2.weight), self._f32("t2b", l2.bias))

    def _project_style(self, s5: torch.Tensor, s6: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.cross_attention2._project_style(s5), self.cross_attention1._project_style(s6)

    def _conv(self, name, x, **kw):
        c = getattr(self, name)
        tr = isinstance(c, nn.ConvTranspose2d)
        wp = self._packed(name, c.weight, x.dtype, tr)
        return ops.conv(x, wp, self._f32(name + ".b", c.bias), ksize=3, stride=c.stride[0], transposed=tr, **kw)

    def _forward_nhwc(self, z: torch.Tensor, temb: torch.Tensor, kv2: torch.Tensor, kv1: torch.Tensor, *,
                      ddim: Optional[ops.Ddim] = None, want_eps: bool = True, eps_out: Optional[torch.Tensor] = None,
                      taps: Optional[dict] = None) -> Optional[torch.Tensor]:
        """z: [B,h,w,Cin] in the activation dtype; temb [B,128] fp32; kv2/kv1: projected style
        maps for cross_attention2 / cross_attention1.  Returns eps_hat [B,h,w,Cout] fp32 (and/or
        runs the fused DDIM update)."""
        z1 = self._conv("enc1", z, act=ACT_RELU)
        z2 = self._conv("enc2", z1, act=ACT_RELU, add_nc=temb)
        z3 = self._conv("enc3", z2, act=ACT_RELU)
        a3 = self.cross_attention2._forward_nhwc(z3, kv2, kv_static=True)      # the projections precede enc1..enc3
        z4 = self._conv("enc4", a3, act=ACT_RELU)
        a4 = self.cross_attention1._forward_nhwc(z4, kv1, kv_static=True)
        bt = self._conv("bottleneck", a4, act=ACT_RELU)
        d4 = self._conv("dec4", bt, act=ACT_RELU, add=z3)
        d3 = self._conv("dec3", d4, act=ACT_RELU, add=z2)
        d2 = self._conv("dec2", d3, act=ACT_RELU, add=z1)
        if taps is not None:
            taps.update(z1=z1, z2=z2, z3=z3, a3=a3, z4=z4, a4=a4, bt=bt, d4=d4, d3=d3, d2=d2)
        return self._conv("dec1", d2, act=ACT_NONE, out_dtype=torch.float32, ddim=ddim, want_y=want_eps, out=eps_out)

    def forward(self, z, t, style_embedding: dict = None):
        _need_cuda(z, "UNet")
        if style_embedding is None:
            raise MstkError("UNet.forward needs style_embedding with 's5' and 's6' (reference model.py:211,214)")
        adt = _act_dtype(self._precision())
        zl = ops.nchw_to_nhwc(z.detach().float().contiguous(), adt)
        nh = getattr(style_embedding, "_nhwc", None)
        if nh and nh["s5"].dtype == adt:
            s5, s6 = nh["s5"], nh["s6"]
        else:
            s5 = ops.nchw_to_nhwc(style_embedding["s5"].detach().float().contiguous(), adt)
            s6 = ops.nchw_to_nhwc(style_embedding["s6"].detach().float().contiguous(), adt)
        if (s5.shape[1] * 4, s5.shape[2] * 4) != tuple(zl.shape[1:3]) or (s6.shape[1] * 8, s6.shape[2] * 8) != tuple(zl.shape[1:3]):
            raise MstkError("UNet: s5 / s6 must be 1/4 and 1/8 of the latent resolution")
        kv2, kv1 = self._project_style(s5, s6)
        eps = self._forward_nhwc(zl, self._time_embedding(t.to(z.device)), kv2, kv1)
        return _guard("UNet", self, ops.nhwc_to_nchw(eps), z, *style_embedding.values())


class LDM(_Kernelised):
    """reference models/model.py:249-559: VAE encoder/decoder + UNet + scheduler + style
    encoder, the training forward and the two deterministic DDIM-style sampling loops.

    Extra (non-reference) attributes, all optional:
      precision            'fp32' | 'bf16' via ``set_precision``
      use_cuda_graphs      capture the whole N-1 step sampling loop in one CUDA graph
      collect_sampling_logs  fill sampling_logs['pred_x0'/'noise_pred'] (the reference always
                           clones them and host-syncs every step, model.py:461-463; off by default
                           because it is pure overhead -- 'timesteps' is always filled)
    """

    def __init__(self, latent_dim, pretrained_path: str = "models/pretrained/", pretraind_filename: str = "ldm.pth",
                 num_timesteps=config["forward_diffusion_num_timesteps"], load_full_model=False):
        super().__init__()
        from .loss import VGGishFeatureLoss
        self.encoder = SpectrogramEncoder(latent_dim=latent_dim)
        self.decoder = SpectrogramDecoder(latent_dim=latent_dim)
        self.unet = UNet(in_channels=latent_dim, out_channels=latent_dim, num_filters=64)
        self.noise_scheduler = ForwardDiffusion(num_timesteps=num_timesteps)
        self.style_encoder = StyleEncoder(in_channels=1, num_filters=64)
        self.num_timesteps = num_timesteps
        self.feature_loss_net = VGGishFeatureLoss()
        self.use_cuda_graphs = True
        self.collect_sampling_logs = False

        if pretrained_path:
            loaded_full = False
            if load_full_model:
                try:
                    self._load_full(pretrained_path + pretraind_filename)
                    loaded_full = True
                except (FileNotFoundError, RuntimeError) as e:   # same fallback as model.py:316-318
                    print(f"Could not load full LDM model: {e}")
                    print("Falling back to loading just encoder/decoder weights")
            if loaded_full:
                # reference returns here (model.py:315): the loaded unet/style encoder are kept
                self.encoder.eval(); self.decoder.train(); self.unet.train(); self.style_encoder.train()
                self._freeze_encoder()
                return
            dev = "cuda" if torch.cuda.is_available() else "cpu"
            self.encoder.load_state_dict(torch.load(pretrained_path + "encoder.pth", map_location=dev))
            self.decoder.load_state_dict(torch.load(pretrained_path + "decoder.pth", map_location=dev))
            print("Loaded pretrained weights from", pretrained_path)
            self._freeze_encoder()
            self.encoder.eval(); self.decoder.train(); self.style_encoder.train()
        # the reference re-creates these three after loading (model.py:350-353): a second RNG draw
        self.unet = UNet(in_channels=latent_dim, out_channels=latent_dim, num_filters=64)
        self.noise_scheduler = ForwardDiffusion(num_timesteps=num_timesteps)
        self.style_encoder = StyleEncoder(in_channels=1, num_filters=64)
        self.num_timesteps = num_timesteps

    def _freeze_encoder(self):
        for p in self.encoder.parameters():
            p.requires_grad = False
        for p in self.decoder.parameters():
            p.requires_grad = True

    def _load_full(self, path: str):
        """prefix-split loader of a full checkpoint (model.py:262-301)"""
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        sd = torch.load(path, map_location=dev)
        for prefix, mod in (("encoder.", self.encoder), ("decoder.", self.decoder), ("unet.", self.unet),
                            ("style_encoder.", self.style_encoder), ("noise_scheduler.", self.noise_scheduler)):
            mod.load_state_dict({k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)})
        print(f"Loaded full pretrained LDM components from {path}")

    # ------------------------------------------------------------------------------
    def _spec_nhwc(self, x: torch.Tensor, who: str) -> torch.Tensor:
        _need_cuda(x, who)
        x = x.detach().float().contiguous()
        if x.dim() != 4 or x.shape[1] != 1:
            raise MstkError(f"{who}: expected a [B,1,H,W] spectrogram, got {tuple(x.shape)}")
        if x.shape[2] % 64 or x.shape[3] % 64:
            # the reference fails later with a skip-connection size mismatch
            raise MstkError(f"{who}: H and W must be multiples of 64 (got {tuple(x.shape[2:])})")
        return x.view(x.shape[0], x.shape[2], x.shape[3], 1)

    def forward(self, x, style, t):
        """model.py:355-379.  Returns dict(z_t, noise, noise_pred, z_0, reconstructed), NCHW fp32.  With gradients
        enabled the outputs carry an autograd node whose backward is the explicit reverse sweep (autograd.py)."""
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            from .autograd import ldm_forward
            return ldm_forward(self, x, style, t)
        adt = _act_dtype(self._precision())
        xs = self._spec_nhwc(x, "LDM.forward")
        ss = self._spec_nhwc(style, "LDM.forward")
        dev = xs.device
        z0 = self.encoder._forward_nhwc(xs)                       # [B,h,w,C] fp32
        styles = self.style_encoder._forward_nhwc(ss)
        b, h, w, c = z0.shape
        t = t.to(device=dev, dtype=torch.int64).contiguous()
        # noise drawn in the reference's NCHW shape so the generator stream matches model.py:110
        noise = torch.randn_like(torch.empty((b, c, h, w), dtype=torch.float32, device=dev))
        noise_l = ops.nchw_to_nhwc(noise, torch.float32)
        z_t, z_t_lp = ops.q_sample(z0, noise_l, self.noise_scheduler.alpha_bar_t, t, lp=(adt == torch.bfloat16))
        kv2, kv1 = self.unet._project_style(styles[4], styles[5])
        eps = self.unet._forward_nhwc(z_t_lp if z_t_lp is not None else z_t, self.unet._time_embedding(t), kv2, kv1)
        z0p, z0p_lp = ops.predict_x0(z_t, eps, self.noise_scheduler.alpha_bar_t, t, lp=(adt == torch.bfloat16))
        rec = self.decoder._forward_nhwc(z0p_lp if z0p_lp is not None else z0p, unit_range=True)
        return {"z_t": ops.nhwc_to_nchw(z_t), "noise": noise, "noise_pred": ops.nhwc_to_nchw(eps),
                "z_0": ops.nhwc_to_nchw(z0), "reconstructed": rec.view(b, 1, rec.shape[1], rec.shape[2])}

    # ------------------------------------------------------------------------------
    # sampling
    # ------------------------------------------------------------------------------
    def _schedule_host(self):
        ab = self.noise_scheduler.alpha_bar_t

        def build():
            a = ab.detach().float().cpu()
            return torch.sqrt(a).tolist(), torch.sqrt(1 - a).tolist()

        return self._cached(("sched_host",), (ab,), build)

    def _step_times(self, times: List[int], b: int, device) -> torch.Tensor:
        """fp32 timestep per (step, sample).  Cached per (schedule, batch, device): the H2D copy must happen outside
        any graph capture (the whole-call graphs look it up while capturing; their eager warm-up fills the cache)."""
        cache = self.__dict__.setdefault("_tt_cache", {})
        key = (tuple(times), int(b), str(device))
        tt = cache.get(key)
        if tt is None:
            if torch.cuda.is_current_stream_capturing():
                raise MstkError("internal: step-time table requested for the first time inside a graph capture")
            tt = torch.tensor(times[:-1], dtype=torch.float32).repeat_interleave(b).to(device)
            if len(cache) >= 16:
                cache.clear()
            cache[key] = tt
        return tt

    def _loop_body(self, x: torch.Tensor, s5: torch.Tensor, s6: torch.Tensor, times: List[int], eta: float, logs,
                   tt: torch.Tensor, kv=None):
        """N-1 UNet evaluations, each ending in the fused DDIM update (in place on ``x``).  ``kv``: the projected
        style maps when the caller has them already (computed on a concurrent branch)."""
        adt = _act_dtype(self._precision())
        b = x.shape[0]
        nsteps = len(times) - 1
        sab, s1m = self._schedule_host()
        if kv is None:
            ops.RECORD_NOTE = "hoisted"
            kv2, kv1 = self.unet._project_style(s5, s6)
            ops.RECORD_NOTE = ""
        else:
            kv2, kv1 = kv
        # every sample of a step shares its timestep: the time MLP runs on the N-1 distinct values, one broadcast
        # kernel lays the rows out per (step, sample) for the conv epilogue's add_nc
        tu = self.unet._time_embedding(tt.view(nsteps, b)[:, 0].contiguous())
        temb = ops.add_bcast(torch.zeros((nsteps, b, tu.shape[1]), dtype=torch.float32, device=x.device), add_nc=tu)
        x_lp = ops.cast(x, adt) if adt != torch.float32 else None
        self._loop_steps(x, x_lp, temb, kv2, kv1, times, eta, logs)
        return x

    def _loop_steps(self, x, x_lp, temb, kv2, kv1, times: List[int], eta: float, logs):
        sab, s1m = self._schedule_host()
        for i in range(len(times) - 1):
            t, tn = times[i], times[i + 1]
            d = ops.Ddim(x_cur=x, x_next=x, x_next_lp=x_lp, sab_t=sab[t], s1m_t=s1m[t], sab_n=sab[tn], s1m_n=s1m[tn],
                         eta=float(eta), x0_pred=logs["pred_x0"][i] if logs is not None else None)
            self.unet._forward_nhwc(x_lp if x_lp is not None else x, temb[i], kv2, kv1, ddim=d, want_eps=logs is not None,
                                    eps_out=logs["noise_pred"][i] if logs is not None else None)

    def _weights_sig(self):
        return tuple((p.data_ptr(), p._version) for p in list(self.unet.parameters()) + [self.noise_scheduler.alpha_bar_t])

    def _weights_sig_full(self):
        mods = (self.encoder, self.decoder, self.unet, self.style_encoder)
        return tuple((p.data_ptr(), p._version) for m in mods for p in m.parameters()) + \
            ((self.noise_scheduler.alpha_bar_t.data_ptr(), self.noise_scheduler.alpha_bar_t._version),)

    def _bn_training(self) -> bool:
        return any(isinstance(m, nn.BatchNorm2d) and m.training for m in list(self.encoder.modules()) + list(self.decoder.modules()))

    def _graph_call(self, key, inputs, body, warm):
        """A whole public call (VAE encode, style encoder, q-sample, the N-1 step loop, decode) as ONE CUDA graph per
        (shapes, schedule, precision, module modes): ``inputs`` are copied into static buffers, the graph is replayed,
        the outputs are cloned.  ``body(*static_inputs)`` is captured; ``warm(*static_inputs)`` is a short eager run that
        fills the packed-weight caches and the split-K workspace first (BatchNorm buffers it moves are restored)."""
        graphs = self.__dict__.setdefault("_graphs", {})
        sig = self._weights_sig_full()
        ent = graphs.get(key)
        if ent is None or ent["sig"] != sig:
            ent = {"sig": sig, "in": tuple(t.clone() for t in inputs)}
            bn_train = self._bn_training()
            bufs = [b for m in (self.encoder, self.decoder) for b in m.buffers()] if bn_train else []
            snap = [b.clone() for b in bufs]
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            self.use_cuda_graphs = False           # the warm-up runs its short loop eagerly
            try:
                with torch.cuda.stream(side):
                    warm(*ent["in"])
            finally:
                self.use_cuda_graphs = True
            torch.cuda.current_stream().wait_stream(side)
            for b, s0 in zip(bufs, snap):          # train-mode BatchNorm: the warm-up must not count as a forward pass
                b.copy_(s0)
            g = torch.cuda.CUDAGraph()
            n0 = ops.LAUNCHES
            with ops.graph_capture(g):
                ent["out"] = body(*ent["in"])
            ent["graph"] = g
            ent["launches"], ops.LAUNCHES = ops.LAUNCHES - n0, n0      # capture enqueues nothing
            # the graph holds raw pointers to the packed weights: keep them alive with it
            ent["keep"] = [dict(m.__dict__.get("_mstk_cache", {})) for m in self.modules()]
            if len(graphs) >= 8:
                graphs.clear()
            graphs[key] = ent
        for dst, src in zip(ent["in"], inputs):
            dst.copy_(src)
        ent["graph"].replay()
        ops._count(ent["launches"])
        if self._bn_training():                    # running buffers moved behind the version counters: eval folds are stale
            for m in (self.encoder, self.decoder):
                cache = m.__dict__.get("_mstk_cache")
                if cache:
                    for k in [k for k in cache if isinstance(k, tuple) and len(k) == 3 and k[1] == "fold"]:
                        del cache[k]
        return tuple(o.clone() for o in ent["out"])

    def _side_streams(self, device):
        ss = self.__dict__.get("_side")
        if ss is None or ss[0].device != torch.device(device):
            ss = (torch.cuda.Stream(device=device), torch.cuda.Stream(device=device))
            self.__dict__["_side"] = ss
        return ss

    def _whole_call_graph_ok(self) -> bool:
        return self.use_cuda_graphs and not torch.cuda.is_current_stream_capturing() and not self.collect_sampling_logs

    def _run_loop(self, z_t: torch.Tensor, s5: torch.Tensor, s6: torch.Tensor, times: List[int], eta: float):
        """z_t [B,h,w,C] fp32 NHWC -> (x [B,h,w,C] fp32, logs-or-None).  Eager, or one CUDA graph
        replay when ``use_cuda_graphs``."""
        nsteps = len(times) - 1
        b = z_t.shape[0]
        logs = None
        if self.collect_sampling_logs and nsteps > 0:
            logs = {"pred_x0": torch.empty((nsteps,) + tuple(z_t.shape), dtype=torch.float32, device=z_t.device),
                    "noise_pred": torch.empty((nsteps,) + tuple(z_t.shape), dtype=torch.float32, device=z_t.device)}
        if nsteps <= 0:
            return z_t.clone(), logs
        if not self.use_cuda_graphs or torch.cuda.is_current_stream_capturing():
            return self._loop_body(z_t.clone(), s5, s6, times, eta, logs, self._step_times(times, b, z_t.device)), logs
        key = (tuple(z_t.shape), tuple(s5.shape), tuple(s6.shape), s5.dtype, tuple(times), float(eta), self._precision(),
               logs is not None)
        graphs = self.__dict__.setdefault("_graphs", {})
        sig = self._weights_sig()
        ent = graphs.get(key)
        if ent is None or ent["sig"] != sig:
            ent = {"sig": sig, "x": z_t.clone(), "s5": s5.clone(), "s6": s6.clone(), "logs": logs,
                   "tt": self._step_times(times, b, z_t.device), "tt2": self._step_times(times[:2], b, z_t.device)}
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):          # eager warm-up: lazy init, packed-weight caches, workspace
                self._loop_body(ent["x"], ent["s5"], ent["s6"], times[:2], eta, None, ent["tt2"])
            torch.cuda.current_stream().wait_stream(side)
            g = torch.cuda.CUDAGraph()
            n0 = ops.LAUNCHES
            with ops.graph_capture(g):
                self._loop_body(ent["x"], ent["s5"], ent["s6"], times, eta, logs, ent["tt"])
            ent["graph"] = g
            ent["launches"], ops.LAUNCHES = ops.LAUNCHES - n0, n0      # capture enqueues nothing
            # the graph holds raw pointers to the packed weights: keep them alive with it
            ent["keep"] = [dict(m.__dict__.get("_mstk_cache", {})) for m in self.unet.modules()]
            if len(graphs) >= 8:
                graphs.clear()
            graphs[key] = ent
        ent["x"].copy_(z_t); ent["s5"].copy_(s5); ent["s6"].copy_(s6)
        ent["graph"].replay()
        ops._count(ent["launches"])
        return ent["x"].clone(), ent["logs"]

    def _sample(self, z_t_nhwc, s5, s6, times, eta):
        x, logs = self._run_loop(z_t_nhwc, s5, s6, times, eta)
        return x, logs

    @staticmethod
    def _times(first: int, n: int) -> List[int]:
        # torch.linspace(first, 0, n).long() as at model.py:420, 514 (fp32 linspace, truncation)
        return torch.linspace(first, 0, n).long().tolist()

    def _style_nhwc(self, style_embedding, adt):
        nh = getattr(style_embedding, "_nhwc", None)
        if nh and nh["s5"].dtype == adt:
            return nh["s5"], nh["s6"]
        return (ops.nchw_to_nhwc(style_embedding["s5"].detach().float().contiguous(), adt),
                ops.nchw_to_nhwc(style_embedding["s6"].detach().float().contiguous(), adt))

    def _public_logs(self, times, logs):
        out = {"timesteps": [int(t) for t in times[:-1]], "pred_x0": [], "noise_pred": []}
        if logs is not None:
            out["pred_x0"] = [ops.nhwc_to_nchw(logs["pred_x0"][i]) for i in range(len(times) - 1)]
            out["noise_pred"] = [ops.nhwc_to_nchw(logs["noise_pred"][i]) for i in range(len(times) - 1)]
        return out

    def _ddim_public(self, z_t, style_embedding, times, eta):
        _need_cuda(z_t, "LDM sampling")
        adt = _act_dtype(self._precision())
        s5, s6 = self._style_nhwc(style_embedding, adt)
        x0 = ops.nchw_to_nhwc(z_t.detach().float().contiguous(), torch.float32)
        x, logs = self._sample(x0, s5, s6, times, eta)
        return ops.nhwc_to_nchw(x), self._public_logs(times, logs)

    def style_conditioned_ddim_sample(self, z_t, style_embedding, timesteps=100, eta=0.0):
        """model.py:409-465: times = linspace(T-1, 0, timesteps).long(); timesteps-1 UNet calls."""
        return self._ddim_public(z_t, style_embedding, self._times(self.num_timesteps - 1, timesteps), eta)

    def content_style_ddim_sample(self, z_t, style_embedding, timesteps=250, eta=0.0):
        """model.py:503-559: times = linspace(timesteps-1, 0, timesteps).long()."""
        return self._ddim_public(z_t, style_embedding, self._times(timesteps - 1, timesteps), eta)

    def style_ddim_sample_wrapper(self, z_shape, style_spec, timesteps=100, eta=0.0):
        """model.py:382-407.  Returns decoded spectrograms in [0,1]."""
        ss = self._spec_nhwc(style_spec, "style_ddim_sample_wrapper")
        z_t = torch.randn(z_shape).to(ss.device).float().contiguous()  # CPU draw then move, as model.py:395

        def body(ss_, z_t_, n=timesteps):
            styles = self.style_encoder._forward_nhwc(ss_)
            x, _ = self._sample(ops.nchw_to_nhwc(z_t_, torch.float32), styles[4], styles[5], self._times(self.num_timesteps - 1, n), eta)
            return (self.decoder._forward_nhwc(x, unit_range=True),)

        if self._whole_call_graph_ok() and timesteps >= 2:
            self._step_times(self._times(self.num_timesteps - 1, timesteps), ss.shape[0], ss.device)
            key = ("style_only", tuple(ss.shape), tuple(z_t.shape), int(timesteps), float(eta), self._precision(), self._bn_training())
            (dec,) = self._graph_call(key, (ss, z_t), body, lambda a, b: body(a, b, n=2))
        else:
            (dec,) = body(ss, z_t)
        return dec.view(dec.shape[0], 1, dec.shape[1], dec.shape[2])

    def transfer_batches(self, batches, num_timesteps=250, eta=0.0, ring: int = 0, depth: int = 2):
        """Serving loop over HOST batches (no reference counterpart: the reference's scripts call the wrapper once per batch and
        copy around it).  ``batches`` yields (content, style) CPU tensors, pinned for asynchronous copies; for each one, in
        order, this yields the wrapper's (decoded, z_t_decoded) as pinned CPU tensors.  The upload of batch i+1 and the
        download of batch i-1 run on copy streams under the kernels of batch i (ops.device_prefetch; the wrapper's outputs
        are clones, so the next replay does not touch what the download reads); each result is handed out ``depth`` calls
        late, after its download event (depth 2: the host may fall ~two calls behind -- a garbage-collection pause, a page
        fault -- before the GPU runs dry; depth 1 is the lowest latency).  ``ring`` = 0: every result lands in freshly allocated pinned tensors; ``ring`` >= 3: results
        cycle through that many pre-allocated pinned pairs (no host allocation inside the loop) and a yielded pair is
        overwritten ``ring - depth - 1`` results later -- for consumers that are done with a result before asking for the next."""
        dev = next(self.parameters()).device
        if dev.type != "cuda":
            raise MstkError("transfer_batches: the model must be on a CUDA device (the kernels have no CPU path)")
        depth = max(1, int(depth))
        if ring and ring < depth + 2:
            raise MstkError(f"transfer_batches: ring must be 0 or at least depth + 2 = {depth + 2} (pairs in flight plus the one with the consumer)")
        down = torch.cuda.Stream(dev)
        pending, slots, n = [], [], 0
        for c, s in ops.device_prefetch(batches, dev):
            dec, ztd = self.content_style_transfer_wrapper(c, s, num_timesteps=num_timesteps, eta=eta)
            main = torch.cuda.current_stream(dev)
            ready = torch.cuda.Event()
            ready.record(main)
            with torch.cuda.stream(down):
                down.wait_event(ready)
                if ring and len(slots) == ring:
                    host = slots[n % ring]
                else:
                    host = tuple(torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in (dec, ztd))
                    if ring:
                        slots.append(host)
                n += 1
                outs = tuple(h.copy_(t, non_blocking=True) for h, t in zip(host, (dec, ztd)))
                done = torch.cuda.Event()
                done.record(down)
            dec.record_stream(down); ztd.record_stream(down)
            pending.append((outs, done))
            if len(pending) > depth:
                first, ev = pending.pop(0)
                ev.synchronize()
                yield first
        for first, ev in pending:
            ev.synchronize()
            yield first

    def content_style_transfer_wrapper(self, content_spec, style_spec, num_timesteps=250, eta=0.0):
        """model.py:468-501.  Returns (decoded in [0,1], z_t_decoded in [-1,1])."""
        cs = self._spec_nhwc(content_spec, "content_style_transfer_wrapper")
        ss = self._spec_nhwc(style_spec, "content_style_transfer_wrapper")
        dev = cs.device
        b, hh, ww = cs.shape[0], cs.shape[1] // 8, cs.shape[2] // 8
        c = self.encoder.encoder[6].out_channels
        # the q-sample noise is drawn with torch's generator as the reference does (model.py:110, the only draw of the call)
        noise = torch.randn_like(torch.empty((b, c, hh, ww), dtype=torch.float32, device=dev))

        def body(cs_, ss_, noise_, n=num_timesteps):
            times = self._times(n - 1, n)
            # Inside the whole-call graph the call forks: the style side (style encoder + composed projections) runs
            # beside the VAE encoder, and the decode of z_t (which the loop does not need) beside the N-1 latency-bound
            # steps.  Not with train-mode BatchNorm: the two decoder passes update the running buffers in call order.
            fork = torch.cuda.is_current_stream_capturing() and not self._bn_training()
            main = torch.cuda.current_stream()
            if fork:
                s_a, s_b = self._side_streams(dev)
                s_a.wait_stream(main)
                with torch.cuda.stream(s_a), ops.workspace_slot(1):
                    styles = self.style_encoder._forward_nhwc(ss_)
                    kv = self.unet._project_style(styles[4], styles[5])
            z0 = self.encoder._forward_nhwc(cs_)
            t = torch.full((b,), n - 1, dtype=torch.int64, device=dev)
            z_t, _ = ops.q_sample(z0, ops.nchw_to_nhwc(noise_, torch.float32), self.noise_scheduler.alpha_bar_t, t)
            if not fork:
                styles = self.style_encoder._forward_nhwc(ss_)
                x, _ = self._sample(z_t, styles[4], styles[5], times, eta)
                return self.decoder._forward_nhwc(x, unit_range=True), self.decoder._forward_nhwc(z_t)
            s_b.wait_stream(main)
            with torch.cuda.stream(s_b), ops.workspace_slot(2):
                ztd_ = self.decoder._forward_nhwc(z_t)
            main.wait_stream(s_a)
            x = self._loop_body(z_t.clone(), styles[4], styles[5], times, eta, None, self._step_times(times, b, dev), kv=kv)
            dec_ = self.decoder._forward_nhwc(x, unit_range=True)
            main.wait_stream(s_b)
            return dec_, ztd_

        if self._whole_call_graph_ok() and num_timesteps >= 2:
            self._step_times(self._times(num_timesteps - 1, num_timesteps), b, dev)
            key = ("transfer", tuple(cs.shape), int(num_timesteps), float(eta), self._precision(), self._bn_training())
            dec, ztd = self._graph_call(key, (cs, ss, noise), body, lambda a, s_, e_: body(a, s_, e_, n=2))
        else:
            dec, ztd = body(cs, ss, noise)
        return dec.view(b, 1, dec.shape[1], dec.shape[2]), ztd.view(b, 1, ztd.shape[1], ztd.shape[2])
