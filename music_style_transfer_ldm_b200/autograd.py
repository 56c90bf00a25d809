"""Autograd face of the drop-in modules: the reference trains with ``loss.backward()`` on module outputs
(models/train.py:176-201), so gradients must flow through ``LDM.forward`` and the loss functions.

There is no autograd tape inside libmstk: every gradient op is a C-ABI kernel.  The nodes here splice the
explicit reverse sweep of the trainer (train.py: ``_forward_keep`` / ``_backward_sweep``) into torch's graph:

  LDMForwardFn   LDM.forward (model.py:355-379) as ONE node: forward keeps the intermediates, backward takes
                 dL/d(noise_pred) and dL/d(reconstructed) and returns the gradients of the decoder, UNet and
                 style-encoder parameters (the encoder is frozen in the LDM stage, model.py:340-341; ``z_t`` and
                 ``z_0`` depend on no trainable parameter and are marked non-differentiable).
  MseFn, KlFn    loss.py's reductions with their analytic gradients (``diffusion_loss``, the MSE and KL terms of
                 ``compression_loss``); the upstream scalar stays on the device (mstk_scale_dev).
  Unsupported    every other module forward taken with gradients enabled returns its output through this node:
                 the forward value is the usual one, a backward pass THROUGH it raises MstkError instead of
                 silently producing no gradient (inputs were detached without a word).

So the reference's training step -- ``out = model(x, style, t)``; the four losses; ``total.backward()``;
``torch.optim.Adam.step()`` -- runs unchanged on this package (tests/test_gpu_autograd.py compares the parameter
gradients with the oracle's autograd).  ``LDMTrainer`` remains the fast path (flat arenas, fused Adam, one graph).
"""
from __future__ import annotations

from typing import List

import torch

from . import model as Mo
from . import ops
from .ops import MstkError


def _trainable(model) -> List[torch.nn.Parameter]:
    from .train import _backward_order
    return _backward_order(model)


class _Sweep:
    """the trainer's two sweep halves without its arenas: gradients accumulate in a fresh flat buffer per backward"""

    def __new__(cls, model):
        from .train import LDMTrainer

        class _S(LDMTrainer):          # borrow the sweep methods; never runs LDMTrainer.__init__
            def __init__(self, m):
                self.model, self.world, self.group, self.grads = m, 1, None, {}

            def _G(self, p):
                return self.grads[id(p)]

            def _sync_sums(self, sums):
                return None

            def _reduce_bucket(self, stage):
                return None

        return _S(model)


class LDMForwardFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, x, style, t, noise, *params):
        xs = model._spec_nhwc(x, "LDM.forward")
        ss = model._spec_nhwc(style, "LDM.forward")
        sw = _Sweep(model)
        with torch.no_grad():
            sv = sw._forward_keep(xs, ss, t.to(device=xs.device, dtype=torch.int64).contiguous(), noise)
            z0, z_t, eps_hat, rec = sv[2], sv[5], sv[29], sv[39]
            b = xs.shape[0]
            outs = (ops.nhwc_to_nchw(z_t), ops.nhwc_to_nchw(eps_hat), ops.nhwc_to_nchw(z0), rec.view(b, 1, rec.shape[1], rec.shape[2]))
        ctx.sw, ctx.sv, ctx.params = sw, sv, params
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(outs[0], outs[2])
        return outs

    @staticmethod
    def backward(ctx, d_zt, d_eps, d_z0, d_rec):
        sw, sv, params = ctx.sw, ctx.sv, ctx.params
        rec, eps_hat, adt = sv[39], sv[29], sv[41]
        dev = rec.device
        offs, total = [], 0
        for p in params:
            offs.append(total)
            total += (p.numel() + 3) // 4 * 4
        flat = torch.zeros(total, dtype=torch.float32, device=dev)
        views = [flat[o:o + p.numel()].view(p.shape) for o, p in zip(offs, params)]
        sw.grads = {id(p): v for p, v in zip(params, views)}
        with torch.no_grad(), Mo.bn_sync_group(None, 1):
            if d_rec is None:
                g_rec = torch.zeros(rec.shape, dtype=torch.float32, device=dev)
            else:
                g_rec = d_rec.detach().float().contiguous().view(rec.shape)
            g_eps = None
            if d_eps is not None:
                g_eps = ops.nchw_to_nhwc(d_eps.detach().float().contiguous(), adt)
            sw._backward_sweep(sv, g_rec, den_scale=0.0, g_eps_direct=g_eps)
        sw.grads = {}
        return (None, None, None, None, None) + tuple(v if p.requires_grad else None for p, v in zip(params, views))


def ldm_forward(model, x, style, t):
    """LDM.forward with gradients: same dict as the no-grad path (model.py:355-379)."""
    params = _trainable(model)
    want = {id(p) for p in model.parameters() if p.requires_grad}
    if want != {id(p) for p in params} or not model.decoder.decoder[1].training:
        # not the reference's LDM-stage configuration (encoder frozen, model.py:340-341; model.train(), train.py:212):
        # the values are computed as usual, a backward pass through them says what is missing
        with torch.no_grad():
            out = model.forward(x, style, t)
        who = ("LDM.forward (gradients need exactly decoder + unet + style_encoder trainable -- model._freeze_encoder() -- "
               "and model.train())")
        return {k: (v if k == "noise" else guard(who, model, v, x, style)) for k, v in out.items()}
    Mo._need_cuda(x, "LDM.forward")
    b, lat = x.shape[0], model.encoder.encoder[6].out_channels
    # the q-sample noise is drawn with torch's generator in the reference's randn_like shape (model.py:110)
    noise = torch.randn_like(torch.empty((b, lat, x.shape[2] // 8, x.shape[3] // 8), dtype=torch.float32, device=x.device))
    z_t, eps_hat, z0, rec = LDMForwardFn.apply(model, x, style, t, noise, *params)
    return {"z_t": z_t, "noise": noise, "noise_pred": eps_hat, "z_0": z0, "reconstructed": rec}


class MseFn(torch.autograd.Function):
    """mean((a - b)^2) (loss.py:36, 49)"""

    @staticmethod
    def forward(ctx, a, b):
        a32, b32 = a.detach().float().contiguous(), b.detach().float().contiguous()
        ctx.save_for_backward(a32, b32)
        return ops.mse(a32, b32).squeeze(0)

    @staticmethod
    def backward(ctx, g):
        a32, b32 = ctx.saved_tensors
        gf = g.detach().float().reshape(1).contiguous()
        ga = gb = None
        if ctx.needs_input_grad[0]:
            ga = ops.scale_dev_(ops.mse_grad(a32, b32, 2.0 / a32.numel(), torch.float32), gf)
        if ctx.needs_input_grad[1]:
            gb = ops.scale_dev_(ops.mse_grad(b32, a32, 2.0 / a32.numel(), torch.float32), gf)
        return ga, gb


class KlFn(torch.autograd.Function):
    """mean(0.5 (z^2 - 1 - log(z^2 + 1e-8))) (loss.py:31-32)"""

    @staticmethod
    def forward(ctx, z):
        z32 = z.detach().float().contiguous()
        ctx.save_for_backward(z32)
        return ops.kl(z32).squeeze(0)

    @staticmethod
    def backward(ctx, g):
        (z32,) = ctx.saved_tensors
        return ops.scale_dev_(ops.kl_grad(z32, None, 1.0 / z32.numel()), g.detach().float().reshape(1).contiguous())


class Unsupported(torch.autograd.Function):
    @staticmethod
    def forward(ctx, who, y, *deps):
        ctx.who = who
        return y.view_as(y)

    @staticmethod
    def backward(ctx, g):
        raise MstkError(f"{ctx.who}: backward through this module's forward is not implemented.  Gradients flow through "
                        "LDM.forward and the functions of loss.py (the reference's training step, train.py:176-201) and "
                        "through LDMTrainer / AutoencoderTrainer; wrap inference calls in torch.no_grad().")


def guard(who: str, module: torch.nn.Module, y, *inputs):
    """Module outputs taken with gradients enabled: same values, but a backward pass through them raises instead of
    silently yielding nothing (the kernels see detached tensors)."""
    if not torch.is_grad_enabled():
        return y
    deps = [p for p in module.parameters() if p.requires_grad] + [i for i in inputs if isinstance(i, torch.Tensor) and i.requires_grad]
    if not deps:
        return y
    if isinstance(y, dict):
        out = type(y)({k: Unsupported.apply(who, v, *deps) for k, v in y.items()})
        if hasattr(y, "_nhwc"):
            out._nhwc = y._nhwc
        return out
    return Unsupported.apply(who, y, *deps)
