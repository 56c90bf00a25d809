"""-m gpu: fp32 mode on the tensor cores (csrc/split_f32.cu + the tcgen05 implicit-GEMM kernel): every fp32
contraction whose shape fits the tcgen05 kernel runs as ONE bf16 contraction over [hi | lo | hi] x [hi | hi | lo]
with fp32 accumulation.  Held to 3e-5 relative (max norm) against a plain PyTorch fp32 reference per op -- the
whole-model fp32 bar is 1e-4 -- with every fused epilogue, and against the CUDA-core fp32 kernel."""
import math

import pytest
import torch
import torch.nn.functional as F

from gpu_util import assert_close, conv_ref, gen, nchw, nhwc

pytestmark = pytest.mark.gpu
DEV = "cuda"

CASES = [
    # n, h, w, cin, cout, k, stride, transposed      (the UNet / VAE / style-encoder / VGGish layer shapes)
    (4, 16, 16, 32, 64, 3, 1, False), (4, 16, 16, 64, 128, 3, 2, False), (3, 8, 8, 128, 256, 3, 2, False),
    (2, 4, 4, 256, 512, 3, 2, False), (2, 2, 2, 512, 512, 3, 1, False), (2, 2, 2, 512, 256, 3, 2, True),
    (2, 8, 8, 128, 64, 3, 2, True), (2, 16, 16, 32, 128, 4, 2, True), (2, 32, 32, 128, 64, 4, 2, True),
    (2, 4, 4, 256, 256, 1, 1, False), (1, 16, 48, 32, 64, 3, 1, False), (2, 32, 32, 64, 128, 3, 1, False),
]


def _ops():
    from music_style_transfer_ldm_b200 import ops
    return ops


def test_split3_is_exact_to_2_pow_minus_17():
    ops = _ops()
    x = torch.randn((3, 5, 7, 64), generator=gen(1)) * torch.logspace(-6, 6, 64)
    a = ops.split3(x.to(DEV)).float().cpu()
    w = ops.split3(x.to(DEV), weights=True).float().cpu()
    hi, lo = a[..., :64], a[..., 64:128]
    assert torch.equal(hi, x.bfloat16().float()) and torch.equal(a[..., 128:], hi)
    assert torch.equal(w[..., :64], hi) and torch.equal(w[..., 64:128], hi) and torch.equal(w[..., 128:], lo)
    assert ((hi + lo - x).abs() <= x.abs() * 2.0 ** -16).all()


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(map(str, c)))
def test_conv_fp32_on_tensor_cores(case):
    ops = _ops()
    n, h, w, cin, cout, k, s, tr = case
    g = gen(hash(case) % 1000)
    x = torch.randn((n, cin, h, w), generator=g)
    wt = torch.randn((cin, cout, k, k) if tr else (cout, cin, k, k), generator=g) / math.sqrt(cin * k * k)
    b = torch.randn((cout,), generator=g)
    ref = conv_ref(x, wt[:, :, 0, 0] if k == 1 else wt, b, k, s, tr)
    wp = ops.pack_weight(wt.to(DEV), torch.float32, transposed=tr)
    xd = nhwc(x).to(DEV)
    n0 = ops.LAUNCHES
    y = ops.conv(xd, wp, b.to(DEV), ksize=k, stride=s, transposed=tr)
    assert ops.LAUNCHES - n0 == 3 and y.dtype == torch.float32       # weight split (once) + activation split + tcgen05 conv
    assert_close(nchw(y), ref, 3e-5, "fp32 conv on tcgen05")
    n0 = ops.LAUNCHES
    y2 = ops.conv(xd, wp, b.to(DEV), ksize=k, stride=s, transposed=tr)
    assert ops.LAUNCHES - n0 == 2                                     # the split weight is cached with the packed weight
    assert_close(y2, y, 1e-6, "repeat (split-K partials meet in any order)")
    y_simt = ops.conv(xd, wp, b.to(DEV), ksize=k, stride=s, transposed=tr, backend=ops.BACKEND_SIMT)
    assert_close(y, y_simt, 3e-5, "vs the CUDA-core fp32 kernel")


def test_epilogues_fp32_on_tensor_cores():
    """bias -> BatchNorm sums -> ReLU -> + t_emb[n, c] -> + skip, and the fused DDIM update"""
    ops = _ops()
    g = gen(9)
    n, h, w, cin, cout = 3, 8, 8, 64, 32
    x = torch.randn((n, cin, h, w), generator=g)
    wt = torch.randn((cout, cin, 3, 3), generator=g) / math.sqrt(cin * 9)
    b = torch.randn((cout,), generator=g)
    add = torch.randn((n, cout, h, w), generator=g)
    anc = torch.randn((n, cout), generator=g)
    pre = F.conv2d(x, wt, b, padding=1)
    ref = F.relu(pre) + anc[:, :, None, None] + add
    wp = ops.pack_weight(wt.to(DEV), torch.float32)
    stats = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    y = ops.conv(nhwc(x).to(DEV), wp, b.to(DEV), ksize=3, act=ops.ACT_RELU, add=nhwc(add).to(DEV), add_nc=anc.to(DEV), stats=stats)
    assert_close(nchw(y), ref, 3e-5, "epilogue")
    assert_close(stats[:cout], pre.double().sum(dim=(0, 2, 3)), 3e-5, "stats sum")
    assert_close(stats[cout:], (pre.double() ** 2).sum(dim=(0, 2, 3)), 3e-5, "stats sumsq")
    # fused DDIM update in place on x_cur
    xc = torch.randn((n, cout, h, w), generator=g)
    sab_t, s1m_t, sab_n, s1m_n, eta = 0.9, math.sqrt(1 - 0.81), 0.95, math.sqrt(1 - 0.9025), 0.3
    x0_ref = (xc - s1m_t * pre) / sab_t
    xn_ref = sab_n * x0_ref + s1m_n * pre + eta * (s1m_n - s1m_t) * pre
    xcd = nhwc(xc).to(DEV)
    x0d = torch.empty_like(xcd)
    d = ops.Ddim(x_cur=xcd, x_next=xcd, sab_t=sab_t, s1m_t=s1m_t, sab_n=sab_n, s1m_n=s1m_n, eta=eta, x0_pred=x0d)
    eps = ops.conv(nhwc(x).to(DEV), wp, b.to(DEV), ksize=3, ddim=d, out_dtype=torch.float32)
    assert_close(nchw(eps), pre, 3e-5, "eps")
    assert_close(nchw(xcd), xn_ref, 3e-5, "x_next")
    assert_close(nchw(x0d), x0_ref, 3e-5, "x0_pred")
