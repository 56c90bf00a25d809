"""-m gpu: the halo-resident persistent tcgen05 kernel (csrc/conv_halo.cu) for the throughput-bound 3x3 stride-1 layers
against a plain PyTorch fp32 reference on the bf16-rounded operands, and against the one-tile kernel it replaces."""
import math
import os
import subprocess
import sys

import pytest
import torch

from gpu_util import assert_close, conv_ref, gen, nchw, nhwc

pytestmark = pytest.mark.gpu
DEV = "cuda"

# n, h, w, cin, cout: all take the halo kernel (>= 2 x 148 items of a 16 x 16 pixel block x 128 output channels)
CASES = [
    (160, 16, 16, 64, 512),      # one block per image, 4 output blocks, single channel block
    (40, 32, 64, 128, 256),      # non-square image, 2 channel blocks
    (10, 64, 64, 64, 512),       # VGGish conv2 geometry
    (150, 16, 32, 192, 128),     # 3 channel blocks (odd), one output block
    (64, 16, 16, 512, 1280),     # 8 channel blocks, 10 output blocks: long K loop, many accumulator hand-overs
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(map(str, c)))
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_conv_halo_vs_torch(case, out_dtype):
    from music_style_transfer_ldm_b200 import ops
    n, h, w, cin, cout = case
    g = gen(hash(case) % 1000)
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cout, cin, 3, 3), generator=g) / math.sqrt(cin * 9)).bfloat16()
    b = torch.randn((cout,), generator=g)
    xd = nhwc(x, torch.bfloat16).to(DEV)
    wp = ops.pack_weight(wt.float().to(DEV), torch.bfloat16)
    bd = b.to(DEV)
    y = ops.conv(xd, wp, bd, ksize=3, act=ops.ACT_RELU, out_dtype=out_dtype, backend=ops.BACKEND_TCGEN05)
    torch.cuda.synchronize()
    ref = torch.relu(torch.nn.functional.conv2d(x.float().to(DEV), wt.float().to(DEV), bd, padding=1))
    assert_close(nchw(y), ref.cpu(), 1e-5 if out_dtype == torch.float32 else 8e-3, "halo vs torch")
    # second launch (fresh barriers / TMEM) and no activation
    y2 = ops.conv(xd, wp, None, ksize=3, act=ops.ACT_NONE, out_dtype=torch.float32, backend=ops.BACKEND_TCGEN05)
    ref2 = torch.nn.functional.conv2d(x.float().to(DEV), wt.float().to(DEV), None, padding=1)
    assert_close(nchw(y2), ref2.cpu(), 1e-5, "halo vs torch (no bias, no act)")


def test_halo_kernel_is_the_one_that_runs_and_matches_the_one_tile_kernel():
    """a child interpreter with MSTK_CONV_HALO=0 gives the one-tile kernel's result for the same seeded operands"""
    code = r'''
import sys, torch, math
sys.path.insert(0, %r)
from music_style_transfer_ldm_b200 import ops
g = torch.Generator().manual_seed(5)
x = torch.randn((96, 32, 32, 64), generator=g).bfloat16().cuda()
wp = (torch.randn((9, 128, 64), generator=g) / 24).bfloat16().cuda()
b = torch.randn((128,), generator=g).cuda()
y = ops.conv(x, wp, b, ksize=3, act=ops.ACT_RELU, out_dtype=torch.float32)
torch.save(y.cpu(), sys.argv[1])
'''
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for flag in ("1", "0"):
        path = f"/tmp/halo_cmp_{flag}.pt"
        r = subprocess.run([sys.executable, "-c", code % root, path], env=dict(os.environ, MSTK_CONV_HALO=flag), capture_output=True,
                           text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(torch.load(path))
    assert_close(outs[0], outs[1], 1e-5, "halo vs one-tile kernel")


def _halo_count(kind):
    import ctypes
    from music_style_transfer_ldm_b200 import _lib
    fn = _lib.load().mstk_debug_halo_count
    fn.restype, fn.argtypes = ctypes.c_longlong, [ctypes.c_int]
    return fn(kind)


# n, h, w, cin: ConvTranspose2d(cin, 64, k4, s2, p1) shapes that take the halo kernel (>= 2 x 148 blocks of 16 x 16 input pixels)
CONVT_CASES = [
    (16, 32, 256, 128),      # config-2 decoder layer (128 -> 64 @ 32 x 256 -> 64 x 512)
    (100, 16, 48, 64),       # one channel block, non-square
    (80, 32, 32, 192),       # three channel blocks
]


@pytest.mark.parametrize("case", CONVT_CASES, ids=lambda c: "x".join(map(str, c)))
def test_convt_halo_vs_torch(case):
    from music_style_transfer_ldm_b200 import ops
    n, h, w, cin = case
    cout = 64
    g = gen(hash(case) % 1000)
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cin, cout, 4, 4), generator=g) / math.sqrt(cin * 4)).bfloat16()
    b = torch.randn((cout,), generator=g)
    xd = nhwc(x, torch.bfloat16).to(DEV)
    wp = ops.pack_weight(wt.float().to(DEV), torch.bfloat16, transposed=True)
    before = _halo_count(2)
    y = ops.conv(xd, wp, b.to(DEV), ksize=4, stride=2, transposed=True, act=ops.ACT_RELU, out_dtype=torch.bfloat16)
    torch.cuda.synchronize()
    assert _halo_count(2) == before + 1, "the conv-transpose did not take the halo kernel"
    assert tuple(y.shape) == (n, 2 * h, 2 * w, cout)
    ref = torch.relu(torch.nn.functional.conv_transpose2d(x.float().to(DEV), wt.float().to(DEV), b.to(DEV), stride=2, padding=1))
    assert_close(nchw(y), ref.cpu(), 8e-3, "convT halo vs torch")
    # no bias / no activation, second launch
    y2 = ops.conv(xd, wp, None, ksize=4, stride=2, transposed=True, act=ops.ACT_NONE, out_dtype=torch.bfloat16)
    ref2 = torch.nn.functional.conv_transpose2d(x.float().to(DEV), wt.float().to(DEV), None, stride=2, padding=1)
    assert_close(nchw(y2), ref2.cpu(), 8e-3, "convT halo vs torch (no bias, no act)")


def test_conv_halo_counter():
    from music_style_transfer_ldm_b200 import ops
    x = torch.randn((640, 16, 16, 64), device=DEV).bfloat16()
    wp = (torch.randn((9, 128, 64), device=DEV) / 24).bfloat16()
    before = _halo_count(1)
    ops.conv(x, wp, None, ksize=3)
    assert _halo_count(1) == before + 1


def test_convt_halo_batchnorm_statistics():
    """train-mode decoder layer: per-channel sum / sum of squares of (conv + bias) gathered in the halo kernel's epilogue"""
    from music_style_transfer_ldm_b200 import ops
    n, h, w, cin, cout = 96, 32, 32, 128, 64
    g = gen(21)
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cin, cout, 4, 4), generator=g) / math.sqrt(cin * 4)).bfloat16()
    b = torch.randn((cout,), generator=g)
    wp = ops.pack_weight(wt.float().to(DEV), torch.bfloat16, transposed=True)
    stats = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    before = _halo_count(2)
    y = ops.conv(nhwc(x, torch.bfloat16).to(DEV), wp, b.to(DEV), ksize=4, stride=2, transposed=True, act=ops.ACT_NONE, stats=stats,
                 out_dtype=torch.bfloat16)
    torch.cuda.synchronize()
    assert _halo_count(2) == before + 1
    pre = torch.nn.functional.conv_transpose2d(x.float().to(DEV), wt.float().to(DEV), b.to(DEV), stride=2, padding=1).double()
    assert_close(nchw(y), pre.float().cpu(), 8e-3, "convT halo (stats variant) output")
    assert_close(stats[:cout], pre.sum(dim=(0, 2, 3)), 1e-4, "stats sum")
    assert_close(stats[cout:], (pre ** 2).sum(dim=(0, 2, 3)), 1e-4, "stats sumsq")


# n, h, w, cin, cout: ConvTranspose2d(cin, cout, k3, s2, p1, output_padding 1) = the data gradient of the stride-2 convolutions
CONVT3_CASES = [
    (64, 32, 32, 128, 64),       # style encoder enc2's data gradient geometry
    (40, 16, 32, 256, 128),      # two 64-channel output blocks, four channel blocks, non-square
    (100, 16, 16, 64, 192),      # three output blocks, one channel block
]


@pytest.mark.parametrize("case", CONVT3_CASES, ids=lambda c: "x".join(map(str, c)))
def test_convt3_halo_vs_torch(case):
    from music_style_transfer_ldm_b200 import ops
    n, h, w, cin, cout = case
    g = gen(hash(case) % 1000)
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cin, cout, 3, 3), generator=g) / math.sqrt(cin * 2.25)).bfloat16()
    b = torch.randn((cout,), generator=g)
    xd = nhwc(x, torch.bfloat16).to(DEV)
    wp = ops.pack_weight(wt.float().to(DEV), torch.bfloat16, transposed=True)
    before = _halo_count(3)
    y = ops.conv(xd, wp, None, ksize=3, stride=2, transposed=True, act=ops.ACT_NONE, out_dtype=torch.bfloat16)
    torch.cuda.synchronize()
    assert _halo_count(3) == before + 1, "the k3 conv-transpose did not take the halo kernel"
    assert tuple(y.shape) == (n, 2 * h, 2 * w, cout)
    ref = torch.nn.functional.conv_transpose2d(x.float().to(DEV), wt.float().to(DEV), None, stride=2, padding=1, output_padding=1)
    assert_close(nchw(y), ref.cpu(), 8e-3, "convT k3 halo vs torch")
    y2 = ops.conv(xd, wp, b.to(DEV), ksize=3, stride=2, transposed=True, act=ops.ACT_RELU, out_dtype=torch.bfloat16)
    ref2 = torch.relu(torch.nn.functional.conv_transpose2d(x.float().to(DEV), wt.float().to(DEV), b.to(DEV), stride=2, padding=1, output_padding=1))
    assert_close(nchw(y2), ref2.cpu(), 8e-3, "convT k3 halo vs torch (bias, ReLU)")


# n, h, w (input), cin, cout: Conv2d(cin, cout, k3, s2, p1) on the halo kernel (a 2x2 stride-1 convolution over the four parity planes)
S2_CASES = [
    (16, 64, 512, 64, 128),      # config-2 encoder layer (64 -> 128 @ 64 x 512 -> 32 x 256)
    (80, 64, 64, 128, 256),      # style encoder enc3 geometry: two channel blocks per plane, two output blocks
    (300, 32, 32, 64, 128),      # one 16 x 16 output block per image
]


@pytest.mark.parametrize("case", S2_CASES, ids=lambda c: "x".join(map(str, c)))
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_conv_s2_halo_vs_torch(case, out_dtype):
    from music_style_transfer_ldm_b200 import ops
    n, h, w, cin, cout = case
    g = gen(hash(case) % 1000)
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cout, cin, 3, 3), generator=g) / math.sqrt(cin * 9)).bfloat16()
    b = torch.randn((cout,), generator=g)
    xd = nhwc(x, torch.bfloat16).to(DEV)
    wp = ops.pack_weight(wt.float().to(DEV), torch.bfloat16)
    before = _halo_count(4)
    import ctypes
    from music_style_transfer_ldm_b200 import _lib
    sw = _lib.load().mstk_debug_halo
    sw.argtypes = [ctypes.c_int]
    sw(9)                                   # the stride-2 form is opt-in (measured neutral against conv_tc, see conv_halo.cu)
    try:
        y = ops.conv(xd, wp, b.to(DEV), ksize=3, stride=2, act=ops.ACT_RELU, out_dtype=out_dtype)
        torch.cuda.synchronize()
    finally:
        sw(1)
    assert _halo_count(4) == before + 1, "the stride-2 convolution did not take the halo kernel"
    assert tuple(y.shape) == (n, h // 2, w // 2, cout)
    ref = torch.relu(torch.nn.functional.conv2d(x.float().to(DEV), wt.float().to(DEV), b.to(DEV), stride=2, padding=1))
    assert_close(nchw(y), ref.cpu(), 1e-5 if out_dtype == torch.float32 else 8e-3, "conv s2 halo vs torch")
