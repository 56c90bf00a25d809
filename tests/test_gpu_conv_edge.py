"""-m gpu: the HBM-streaming kernels for Cin == 1 (conv k3) and Cout == 1 (convT k4) layers
(csrc/conv_edge.cu, picked by BACKEND_AUTO) against a plain PyTorch fp32 reference."""
import math

import pytest
import torch
import torch.nn.functional as F

from gpu_util import assert_close, gen, nchw, nhwc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ops():
    from music_style_transfer_ldm_b200 import ops
    return ops


@pytest.mark.parametrize("n,h,w,cout,stride,odt", [(3, 128, 128, 64, 2, torch.bfloat16), (2, 64, 48, 64, 1, torch.float32),
                                                    (2, 30, 22, 24, 2, torch.float32), (1, 128, 1024, 64, 2, torch.bfloat16)])
def test_conv_cin1(n, h, w, cout, stride, odt):
    ops = _ops()
    g = gen(n * 7 + cout)
    x = torch.rand((n, 1, h, w), generator=g)
    wt = torch.randn((cout, 1, 3, 3), generator=g) / 3
    b = torch.randn((cout,), generator=g) * 0.1
    pre = F.conv2d(x, wt, b, stride=stride, padding=1)
    stats = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    wp = ops.pack_weight(wt.to(DEV), torch.float32)
    y = ops.conv(x.view(n, h, w, 1).to(DEV), wp, b.to(DEV), ksize=3, stride=stride, act=ops.ACT_RELU, stats=stats, out_dtype=odt)
    assert y.dtype == odt
    assert_close(nchw(y), F.relu(pre), 1e-5 if odt == torch.float32 else 8e-3, "cin1 conv")
    assert_close(stats[:cout], pre.double().sum(dim=(0, 2, 3)), 1e-5, "stats sum")
    assert_close(stats[cout:], (pre.double() ** 2).sum(dim=(0, 2, 3)), 1e-5, "stats sumsq")
    # against the generic CUDA-core kernel as well
    y2 = ops.conv(x.view(n, h, w, 1).to(DEV), wp, b.to(DEV), ksize=3, stride=stride, act=ops.ACT_RELU, out_dtype=torch.float32,
                  backend=ops.BACKEND_SIMT)
    assert_close(nchw(y2), F.relu(pre), 1e-5, "generic kernel")


@pytest.mark.parametrize("n,h,w,cin,dt", [(3, 64, 64, 64, torch.bfloat16), (2, 16, 24, 64, torch.float32), (1, 5, 7, 8, torch.float32),
                                          (1, 64, 512, 64, torch.bfloat16), (2, 5, 7, 64, torch.bfloat16), (1, 13, 70, 64, torch.bfloat16),
                                          (1, 9, 33, 32, torch.bfloat16)])     # bf16 Cin=64: the mma + col2im kernel, ragged tiles too
def test_convt4_cout1(n, h, w, cin, dt):
    ops = _ops()
    g = gen(n * 11 + cin)
    x = torch.randn((n, cin, h, w), generator=g).to(dt)
    wt = (torch.randn((cin, 1, 4, 4), generator=g) / math.sqrt(cin * 4)).to(dt)
    b = torch.randn((1,), generator=g) * 0.1
    ref = torch.tanh(F.conv_transpose2d(x.float(), wt.float(), b, stride=2, padding=1)) * 0.5 + 0.5
    wp = ops.pack_weight(wt.float().to(DEV), dt, transposed=True)
    y = ops.conv(nhwc(x, dt).to(DEV), wp, b.to(DEV), ksize=4, stride=2, transposed=True, act=ops.ACT_TANH, out_dtype=torch.float32,
                 out_scale=0.5, out_shift=0.5)
    assert tuple(y.shape) == (n, 2 * h, 2 * w, 1)
    assert_close(nchw(y), ref, 2e-5, "convT k4 cout1")


@pytest.mark.parametrize("n,h,w", [(2, 16, 32), (1, 128, 128), (3, 6, 64)])
def test_conv_cin1_k4s2_mma(n, h, w):
    """Conv2d(1, 64, 4, 2, 1), fp32 in -> bf16 out (data gradient of the decoder's last layer): mma.sync kernel with hi/lo
    bf16 operand split, against F.conv2d; tolerance = bf16 rounding of the output"""
    ops = _ops()
    g = gen(n * 3 + h)
    x = torch.randn((n, 1, h, w), generator=g)
    wt = torch.randn((64, 1, 4, 4), generator=g) / 4
    b = torch.randn((64,), generator=g) * 0.1
    ref = F.conv2d(x, wt, b, stride=2, padding=1)
    wp = ops.pack_weight(wt.to(DEV), torch.float32)
    y = ops.conv(nhwc(x, torch.float32).to(DEV), wp, b.to(DEV), ksize=4, stride=2, out_dtype=torch.bfloat16)
    assert tuple(y.shape) == (n, h // 2, w // 2, 64)
    assert_close(nchw(y.float()), ref, 8e-3, "conv k4 s2 cin1 (mma)")
    y0 = ops.conv(nhwc(x, torch.float32).to(DEV), wp, None, ksize=4, stride=2, out_dtype=torch.float32, backend=ops.BACKEND_SIMT)
    yb = ops.conv(nhwc(x, torch.float32).to(DEV), wp, None, ksize=4, stride=2, out_dtype=torch.bfloat16)
    assert (yb.float() - y0.bfloat16().float()).abs().max().item() <= 2e-2 * y0.abs().max().item() / 2      # within an ulp or two of the fp32 path
