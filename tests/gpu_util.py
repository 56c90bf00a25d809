"""Helpers shared by the -m gpu parity tests (CUDA path vs CPU oracle / torch fp32 reference)."""
from __future__ import annotations

import torch
import torch.nn.functional as F

# Tolerances (BASELINE.json north_star): fp32 mode 1e-4 relative, bf16 mode 2e-2 relative.
#   fp32: relative error in the MAX norm, max|got - ref| / max|ref| <= 1e-4 (every element).
#   bf16: relative error in the L2 norm, ||got - ref|| / ||ref|| <= 2e-2, against the FP32
#         reference (stricter than against a bf16-autocast reference, whose own rounding noise
#         is of the same size), plus a max-norm guard of 8e-2 so that no single element is off
#         by more than a few bf16 roundings accumulated through the BatchNorm'd decoder.
TOL = {"fp32": 1e-4, "bf16": 2e-2}
BF16_MAX_GUARD = 8e-2


def rel_err(got: torch.Tensor, ref: torch.Tensor) -> float:
    got = got.detach().double().cpu()
    ref = ref.detach().double().cpu() if isinstance(ref, torch.Tensor) else torch.as_tensor(ref).double()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    den = ref.abs().max().item()
    return (got - ref).abs().max().item() / (den if den > 0 else 1.0)


def rel_l2(got: torch.Tensor, ref) -> float:
    got = got.detach().double().cpu()
    ref = ref.detach().double().cpu() if isinstance(ref, torch.Tensor) else torch.as_tensor(ref).double()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    den = ref.norm().item()
    return (got - ref).norm().item() / (den if den > 0 else 1.0)


# every whole-model comparison is recorded here; tests/conftest.py prints the table in the terminal summary
# (also with -q), so a test log shows BOTH norms of every case, not only pass / fail
REPORT = []


def assert_close(got, ref, tol, what=""):
    e = rel_err(got, ref)
    assert e <= tol, f"{what}: relative error {e:.3e} > {tol:.1e}"


def assert_close_mode(got, ref, precision, what="", max_guard=BF16_MAX_GUARD):
    """the parity bar of a whole-model output in the given precision mode (see TOL above); ``max_guard`` replaces the bf16
    max-norm guard where a case is known to sit next to it (the bar itself, the L2 norm, is never loosened)"""
    e2, em = rel_l2(got, ref), rel_err(got, ref)
    REPORT.append((what, precision, e2, em))
    if precision == "fp32":
        assert em <= TOL["fp32"], f"{what}: relative error {em:.3e} > {TOL['fp32']:.1e}"
        return
    assert e2 <= TOL["bf16"], f"{what}: relative L2 error {e2:.3e} > {TOL['bf16']:.1e} (max-norm {em:.3e})"
    assert em <= max_guard, f"{what}: max-norm relative error {em:.3e} > {max_guard:.1e}"


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
