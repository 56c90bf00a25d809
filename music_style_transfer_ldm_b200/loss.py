"""Drop-in surface of the reference's models/loss.py on libmstk kernels.

  diffusion_loss          loss.py:48-49     one fused squared-difference reduction
  kl_regularization_loss  loss.py:31-32     one fused reduction
  compression_loss        loss.py:34-45     MSE + 0.1 * perceptual + 0.01 * KL
  VGGishFeatureLoss       loss.py:52-101    frozen conv stack (thirdparty/torchvggish/
                                            vggish.py:108-118) + per-sample std-normalised MSE taps
  style_loss              loss.py:104-106

LPIPS (loss.py:6-21) needs the lpips package and AlexNet weights: unavailable offline,
parity unpinned, not provided.  The VGGish weights are likewise not downloadable; the
feature network is built with the reference ARCHITECTURE and default-initialised weights
(load a state_dict to use real ones).  All scalars are returned as 0-dim fp32 CUDA tensors
without host synchronisation.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops
from .config import config
from .ops import ACT_RELU, MstkError

_VGGISH_CFG = [64, "M", 128, "M", 256, 256, "M", 512, 512, "M"]


def _vggish_features() -> nn.Sequential:
    layers, cin = [], 1
    for v in _VGGISH_CFG:
        if v == "M":
            layers.append(nn.MaxPool2d(kernel_size=2, stride=2))
        else:
            layers += [nn.Conv2d(cin, v, kernel_size=3, padding=1), nn.ReLU(inplace=True)]
            cin = v
    return nn.Sequential(*layers)


def _as_cuda_f32(x: torch.Tensor, who: str) -> torch.Tensor:
    if not x.is_cuda:
        raise MstkError(f"{who}: input is on {x.device}; the loss kernels are CUDA (sm_100a) only")
    return x.detach().float().contiguous()


def _mse(a, b, who: str):
    """mean((a - b)^2); with gradients enabled and a differentiable operand, an autograd node (autograd.MseFn)"""
    if torch.is_grad_enabled() and (a.requires_grad or b.requires_grad):
        from .autograd import MseFn
        if not (a.is_cuda and b.is_cuda):
            raise MstkError(f"{who}: inputs must be CUDA tensors; the loss kernels are CUDA (sm_100a) only")
        return MseFn.apply(a, b)
    return ops.mse(_as_cuda_f32(a, who), _as_cuda_f32(b, who)).squeeze(0)


def diffusion_loss(noise_pred, noise_target):
    return _mse(noise_pred, noise_target, "diffusion_loss")


def kl_regularization_loss(latent):
    if torch.is_grad_enabled() and latent.requires_grad:
        from .autograd import KlFn
        if not latent.is_cuda:
            raise MstkError("kl_regularization_loss: input must be a CUDA tensor")
        return KlFn.apply(latent)
    return ops.kl(_as_cuda_f32(latent, "kl_regularization_loss")).squeeze(0)


class VGGishFeatureLoss(nn.Module):
    """loss.py:52-101.  ``features`` has the reference's layer indices (conv at 0,3,6,8,11,13)
    so a reference ``feature_loss_net.features.*`` state_dict loads unchanged.  Always frozen
    and evaluated without gradient, as in the reference (``torch.no_grad()``, loss.py:78)."""

    def __init__(self):
        super().__init__()
        self.features = _vggish_features()
        self.features.eval()
        for p in self.features.parameters():
            p.requires_grad = False
        self.precision = None
        self._cache = {}

    def _packed(self, idx: int, dtype):
        conv = self.features[idx]
        sig = (conv.weight.data_ptr(), conv.weight._version, dtype)
        hit = self._cache.get(idx)
        if hit is None or hit[0] != sig:
            hit = (sig, ops.pack_weight(conv.weight, dtype), conv.bias.detach().float().contiguous())
            self._cache[idx] = hit
        return hit[1], hit[2]

    def _taps(self, x_nhwc_f32: torch.Tensor):
        from .model import _DEFAULT_PRECISION, _act_dtype
        adt = _act_dtype(self.precision or _DEFAULT_PRECISION)
        taps, h = [], x_nhwc_f32
        for idx, layer in enumerate(self.features):
            if isinstance(layer, nn.Conv2d):
                first = layer.in_channels == 1
                wp, b = self._packed(idx, torch.float32 if first else adt)
                h = ops.conv(h, wp, b, ksize=3, stride=1, act=ACT_RELU, out_dtype=adt)
                taps.append(h)
            elif isinstance(layer, nn.MaxPool2d):
                h = ops.maxpool2x2(h)
        return taps

    def _loss3(self, three_nhwc_f32: torch.Tensor, b: int, out_a: torch.Tensor, out_c: torch.Tensor) -> None:
        """Training-step form (train.py:182-186 evaluates the loss on (x, recon) and on (recon, style)): ``three`` =
        [x; recon; style] stacked, one pass of the frozen stack; every tap is ONE streaming kernel that yields both
        pair terms and the pooled input of the next conv."""
        from .model import _DEFAULT_PRECISION, _act_dtype
        adt = _act_dtype(self.precision or _DEFAULT_PRECISION)
        layers = list(self.features)
        convs = [i for i, l in enumerate(layers) if isinstance(l, nn.Conv2d)]
        h = three_nhwc_f32
        for ci, idx in enumerate(convs):
            layer = layers[idx]
            first = layer.in_channels == 1
            if (first and ci == 0 and layer.out_channels == 64 and h.shape[1] % 2 == 0 and h.shape[2] % 2 == 0
                    and idx + 2 < len(layers) and isinstance(layers[idx + 2], nn.MaxPool2d)):
                # the first tap is never materialised: conv(1 -> 64) + ReLU recomputed inside the statistics pass
                wp, bias = self._packed(idx, torch.float32)
                h = ops.feature_first3(h, wp, bias, b, 1.0 / len(convs), out_a, out_c, adt)
                continue
            wp, bias = self._packed(idx, torch.float32 if first else adt)
            h = ops.conv(h, wp, bias, ksize=3, stride=1, act=ACT_RELU, out_dtype=adt)
            last = ci == len(convs) - 1
            pool = (not last) and idx + 2 < len(layers) and isinstance(layers[idx + 2], nn.MaxPool2d)
            vec = 8 if adt == torch.bfloat16 else 4
            if h.shape[3] % vec == 0 and h.shape[1] % 2 == 0 and h.shape[2] % 2 == 0:
                pooled = ops.feature_fused3(h, b, 1.0 / len(convs), out_a, out_c, pool)
            else:       # odd shapes: the separate kernels
                ops.feature_mse3(h, b, 1.0 / len(convs), out_a, out_c)
                pooled = ops.maxpool2x2(h) if pool else None
            if pool:
                h = pooled

    def forward(self, predicted, target):
        p = _as_cuda_f32(predicted, "VGGishFeatureLoss")
        t = _as_cuda_f32(target, "VGGishFeatureLoss")
        if p.shape != t.shape or p.shape[1] != 1:
            raise MstkError("VGGishFeatureLoss expects two [B,1,H,W] tensors of the same shape")
        b = p.shape[0]
        # one pass over the stack for both images: rows [0,b) = predicted, [b,2b) = target
        both = torch.cat((p, t), dim=0).view(2 * b, p.shape[2], p.shape[3], 1)
        taps = self._taps(both)
        out = torch.zeros(1, dtype=torch.float32, device=p.device)
        for f in taps:
            ops.feature_std_mse(f[:b], f[b:], 1.0 / len(taps), out)
        return out.squeeze(0)


def perceptual_loss(original, reconstructed, feature_extractor_type: str = "vggish", feature_extractor=None):
    if feature_extractor_type == "vggish":
        assert feature_extractor is not None, "Feature extractor must be provided for VGGish"
        return feature_extractor(original, reconstructed)
    raise MstkError("perceptual_loss: the LPIPS branch (reference loss.py:6-21) needs the lpips package and AlexNet "
                    "weights, which are unavailable offline; use feature_extractor_type='vggish'")


def compression_loss(original, reconstructed, latent, feature_extractor):
    mse_loss = _mse(reconstructed, original, "compression_loss")
    perc = perceptual_loss(original, reconstructed, config["compression_feature_extractor"], feature_extractor=feature_extractor)
    return mse_loss + 0.1 * perc + 0.01 * kl_regularization_loss(latent)


def style_loss(reconstructed, style_spec, feature_loss_net):
    return feature_loss_net(reconstructed, style_spec)
