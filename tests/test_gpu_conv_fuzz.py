"""-m gpu: seeded random shapes through the public conv / wgrad entry points with the AUTO backend, against plain PyTorch fp32
on the CPU.  The layer-shape tests (test_gpu_conv_tc / _halo / _edge / _wgrad_tc) pin the kernels the models use; this file walks
the DISPATCH instead: odd heights and widths, batches that do not fill a tile, channel counts on both sides of every kernel's
eligibility rule (Cin 1 / 32-multiples / others, Cout 1 / 16- / 64- / 128-multiples / others), each supported (k, stride,
transposed) form (conv_plan.cu:48-67), both precision modes, and the fused epilogue (bias, activation, per-sample and
full residual adds, channel statistics) in random combinations.  Whatever kernel AUTO picks must give the reference's numbers
(models/model.py builds its layers from exactly these nn.Conv2d / nn.ConvTranspose2d forms)."""
import math
import random

import pytest
import torch
import torch.nn.functional as F

from gpu_util import conv_ref, gen, nchw, nhwc, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"

CIN = [1, 2, 3, 8, 16, 32, 48, 64, 96, 128, 160, 256]
COUT = [1, 2, 5, 16, 32, 64, 96, 128, 192, 256]
FORMS = [(1, 1, False), (3, 1, False), (3, 2, False), (4, 2, False), (3, 2, True), (4, 2, True)]


def _draw(rng: random.Random):
    k, s, tr = rng.choice(FORMS)
    n = rng.choice([1, 1, 2, 3, 5])
    h, w = rng.randint(1, 37), rng.randint(1, 37)
    if k == 4 and not tr:
        h, w = max(h, 2), max(w, 2)
    if rng.random() < 0.3:                       # tile-aligned maps take the tensor-core kernels' fast paths
        h, w = rng.choice([8, 16, 32]), rng.choice([8, 16, 32, 64])
    cin, cout = rng.choice(CIN), rng.choice(COUT)
    if rng.random() < 0.35:                      # pairs the halo / tcgen05 kernels accept
        cin, cout = rng.choice([32, 64, 128]), rng.choice([32, 64, 128, 256])
    return n, h, w, cin, cout, k, s, tr


def _cases(seed, count):
    rng = random.Random(seed)
    return [_draw(rng) + (rng.randrange(1 << 30),) for _ in range(count)]


def _act_ref(ops, pre, act):
    return {ops.ACT_NONE: pre, ops.ACT_RELU: F.relu(pre), ops.ACT_TANH: torch.tanh(pre)}[act]


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
@pytest.mark.parametrize("case", _cases(2026, 48), ids=lambda c: "x".join(map(str, c[:8])))
def test_conv_auto_backend_random_shapes(case, precision):
    from music_style_transfer_ldm_b200 import ops
    n, h, w, cin, cout, k, s, tr, seed = case
    rng, g = random.Random(seed), gen(seed % 100003)
    dt = torch.bfloat16 if precision == "bf16" else torch.float32
    x = torch.randn((n, cin, h, w), generator=g).to(dt)
    wt = (torch.randn((cin, cout, k, k) if tr else (cout, cin, k, k), generator=g) / math.sqrt(cin * k * k)).to(dt)
    bias = torch.randn((cout,), generator=g) if rng.random() < 0.7 else None
    act = rng.choice([ops.ACT_NONE, ops.ACT_RELU, ops.ACT_TANH])
    pre = conv_ref(x.float(), wt.float()[:, :, 0, 0] if k == 1 else wt.float(), bias, k, s, tr)
    ref = _act_ref(ops, pre, act)
    add = torch.randn(pre.shape, generator=g).to(dt) if rng.random() < 0.3 else None
    add_nc = torch.randn((n, cout), generator=g) if rng.random() < 0.3 else None
    if add_nc is not None:
        ref = ref + add_nc[:, :, None, None]
    if add is not None:
        ref = ref + add.float()
    stats = torch.zeros(2 * cout, dtype=torch.float64, device=DEV) if rng.random() < 0.3 else None
    out_dtype = torch.float32 if precision == "fp32" or rng.random() < 0.5 else torch.bfloat16
    wp = ops.pack_weight(wt.float().to(DEV), dt, transposed=tr)
    y = ops.conv(nhwc(x, dt).to(DEV), wp, bias.to(DEV) if bias is not None else None, ksize=k, stride=s, transposed=tr, act=act,
                 add=nhwc(add, dt).to(DEV) if add is not None else None, add_nc=add_nc.to(DEV) if add_nc is not None else None,
                 stats=stats, out_dtype=out_dtype)
    torch.cuda.synchronize()
    assert y.dtype == out_dtype and tuple(y.shape) == (n, ref.shape[2], ref.shape[3], cout)
    # fp32 mode: BASELINE's 1e-4 on the max norm; bf16 operands are exact in the reference, so an fp32 output differs only by
    # accumulation order, a bf16 output by its own rounding (2^-8 relative per element)
    tol = 1e-4 if precision == "fp32" else (2e-5 if out_dtype == torch.float32 else 8e-3)
    e = rel_err(nchw(y), ref)
    assert e <= tol, f"{case[:8]} {precision} out={out_dtype} act={act} add={add is not None} add_nc={add_nc is not None}: {e:.3e} > {tol:.1e}"
    if stats is not None:
        for name, got, want in (("sum", stats[:cout], pre.double().sum(dim=(0, 2, 3))), ("sumsq", stats[cout:], (pre.double() ** 2).sum(dim=(0, 2, 3)))):
            es = rel_err(got, want)
            assert es <= 2e-4, f"{case[:8]} {precision}: channel {name} off by {es:.3e}"
    # the split-K scratch is shared by every later launch of the process: it must have been left zeroed
    assert int(torch.count_nonzero(ops.workspace(torch.device(DEV)))) == 0, f"{case[:8]}: split-K workspace not left zeroed"


@pytest.mark.parametrize("case", _cases(77, 32), ids=lambda c: "x".join(map(str, c[:8])))
def test_wgrad_random_shapes(case):
    """mstk_conv2d_wgrad (tcgen05, edge mma.sync and CUDA-core kernels behind one entry point) against torch autograd"""
    from music_style_transfer_ldm_b200 import ops
    n, h, w, cin, cout, k, s, tr, seed = case
    g = gen(seed % 100003)
    dt = torch.bfloat16
    x = torch.randn((n, cin, h, w), generator=g).to(dt).float()
    wt = (torch.randn((cin, cout, k, k) if tr else (cout, cin, k, k), generator=g) / math.sqrt(cin * k * k)).requires_grad_(True)
    y = conv_ref(x, wt[:, :, 0, 0] if k == 1 else wt, None, k, s, tr)
    dy = torch.randn(y.shape, generator=g).to(dt).float()
    y.backward(dy)
    dw = torch.zeros(wt.shape, device=DEV)
    ops.wgrad(nhwc(x, dt).to(DEV), nhwc(dy, dt).to(DEV), dw, ksize=k, stride=s, transposed=tr)
    torch.cuda.synchronize()
    e = rel_err(dw.cpu(), wt.grad)
    assert e <= 2e-4, f"{case[:8]}: weight gradient off by {e:.3e}"
