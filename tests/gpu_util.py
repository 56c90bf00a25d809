"""Helpers shared by the -m gpu parity tests (CUDA path vs CPU oracle / torch fp32 reference)."""
from __future__ import annotations

import torch
import torch.nn.functional as F

# Tolerances (BASELINE.json north_star): fp32 mode 1e-4 relative, bf16 mode 2e-2 relative.
# "relative" is measured in the max norm:  max|got - ref| / max|ref|.
TOL = {"fp32": 1e-4, "bf16": 2e-2}


def rel_err(got: torch.Tensor, ref: torch.Tensor) -> float:
    got = got.detach().double().cpu()
    ref = ref.detach().double().cpu() if isinstance(ref, torch.Tensor) else torch.as_tensor(ref).double()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    den = ref.abs().max().item()
    return (got - ref).abs().max().item() / (den if den > 0 else 1.0)


def assert_close(got, ref, tol, what=""):
    e = rel_err(got, ref)
    assert e <= tol, f"{what}: relative error {e:.3e} > {tol:.1e}"


def nhwc(x_nchw: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    return x_nchw.permute(0, 2, 3, 1).contiguous().to(dtype)


def nchw(x_nhwc: torch.Tensor) -> torch.Tensor:
    return x_nhwc.float().permute(0, 3, 1, 2).contiguous()


def pack_ref(w: torch.Tensor, transposed: bool) -> torch.Tensor:
    """[k*k][cout][cin] packing, restated with torch for checking mstk_pack_*."""
    if w.dim() == 2:
        return w[None].contiguous()
    if transposed:   # [ci, co, k, k]
        return w.permute(2, 3, 1, 0).reshape(-1, w.shape[1], w.shape[0]).contiguous()
    return w.permute(2, 3, 0, 1).reshape(-1, w.shape[0], w.shape[1]).contiguous()


def conv_ref(x_nchw, w, b, ksize, stride, transposed):
    """plain PyTorch fp32 reference of the contraction (CPU)"""
    if transposed:
        op = 1 if ksize == 3 else 0
        return F.conv_transpose2d(x_nchw, w, b, stride=2, padding=1, output_padding=op)
    if ksize == 1:
        return F.conv2d(x_nchw, w[:, :, None, None] if w.dim() == 2 else w, b)
    return F.conv2d(x_nchw, w, b, stride=stride, padding=1)


def gen(seed):
    return torch.Generator().manual_seed(seed)
