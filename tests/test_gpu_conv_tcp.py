"""-m gpu: the (opt-in) persistent tile loop of the tcgen05 convolution (csrc/conv_tcp_body.cuh: one CTA per SM, smem ring running
across tiles, double-buffered TMEM accumulators) against the one-tile-per-CTA kernel on the same operands (must agree to
fp32 summation order: same MMAs, same K order) and against a plain PyTorch fp32 reference on the bf16-rounded operands.
Shapes: throughput-bound layers of the VGGish stack / VAE / style encoder (several tiles per SM), every epilogue the
persistent kernel carries (bias, ReLU, tanh + affine, BatchNorm sums, fp32 and bf16 output), ragged batch tails."""
import ctypes as C
import math

import pytest
import torch
import torch.nn.functional as F

from gpu_util import assert_close, conv_ref, gen, nchw, nhwc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _lib():
    from music_style_transfer_ldm_b200 import _lib
    lib = _lib.load()
    lib.mstk_debug_tc_persist.argtypes = [C.c_int]
    lib.mstk_debug_tc_last_config.restype = C.POINTER(C.c_int)
    return lib


@pytest.fixture(autouse=True)
def _no_halo_kernel():
    """these tests compare the two tile loops of conv_tc.cu: keep the halo-resident kernel (conv_halo.cu, which takes the
    VGGish-sized and config-2 decoder shapes by default) out of the way"""
    lib = _lib()
    lib.mstk_debug_halo.argtypes = [C.c_int]
    lib.mstk_debug_halo(0)
    yield
    lib.mstk_debug_halo(1)


CASES = [
    # n, h, w, cin, cout, k, stride, transposed
    (16, 32, 256, 128, 64, 4, 2, True),      # config 2 decoder: 4 phases x 1024 tiles
    (16, 16, 128, 32, 128, 4, 2, True),      # config 2 decoder first layer (KC = 32)
    (16, 64, 512, 64, 128, 3, 2, False),     # config 2 encoder
    (48, 64, 64, 64, 128, 3, 1, False),      # VGGish conv2
    (59, 32, 32, 128, 256, 3, 1, False),     # VGGish conv3: one 256-column tile (two column blocks); 59 images: ragged tail
    (117, 16, 16, 256, 512, 3, 1, False),    # VGGish conv5: K = 2304, 4 channel tiles per pixel tile, ragged tail
    (70, 16, 16, 64, 32, 3, 2, True),        # convT k3: phases with 1 / 2 / 2 / 4 taps (different K per tile), BN = 32
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(map(str, c)))
def test_persistent_matches_one_tile_kernel(case):
    from music_style_transfer_ldm_b200 import ops
    lib = _lib()
    n, h, w, cin, cout, k, s, tr = case
    g = gen(hash(case) % 1000)
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cin, cout, k, k) if tr else (cout, cin, k, k), generator=g) / math.sqrt(cin * k * k)).bfloat16().float()
    b = torch.randn((cout,), generator=g)
    xd = nhwc(x.float(), torch.bfloat16).to(DEV)
    wp = ops.pack_weight(wt.to(DEV), torch.bfloat16, transposed=tr)
    bd = b.to(DEV)
    stats_p = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    stats_o = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    try:
        lib.mstk_debug_tc_persist(1)
        y_p = ops.conv(xd, wp, bd, ksize=k, stride=s, transposed=tr, act=ops.ACT_RELU, backend=ops.BACKEND_TCGEN05)
        assert lib.mstk_debug_tc_last_config()[1] == -1, "expected the persistent kernel for this shape"
        y_ps = ops.conv(xd, wp, bd, ksize=k, stride=s, transposed=tr, stats=stats_p, out_dtype=torch.float32, backend=ops.BACKEND_TCGEN05)
        y_pt = ops.conv(xd, wp, bd, ksize=k, stride=s, transposed=tr, act=ops.ACT_TANH, out_scale=0.5, out_shift=0.5, out_dtype=torch.float32,
                        backend=ops.BACKEND_TCGEN05)
        lib.mstk_debug_tc_persist(0)
        y_o = ops.conv(xd, wp, bd, ksize=k, stride=s, transposed=tr, act=ops.ACT_RELU, backend=ops.BACKEND_TCGEN05)
        assert lib.mstk_debug_tc_last_config()[1] >= 1
        y_os = ops.conv(xd, wp, bd, ksize=k, stride=s, transposed=tr, stats=stats_o, out_dtype=torch.float32, backend=ops.BACKEND_TCGEN05)
    finally:
        lib.mstk_debug_tc_persist(0)         # the library default (see conv_tc.cu: measured slower than two resident one-tile CTAs)
    assert torch.equal(y_p, y_o), "bf16 output of the two kernels differs"
    assert torch.equal(y_ps, y_os)
    assert_close(stats_p, stats_o, 1e-6, "BatchNorm sums (fp32 per-CTA partials meet in any order)")
    ref = conv_ref(x.float(), wt[:, :, 0, 0] if k == 1 else wt, b, k, s, tr)
    assert_close(nchw(y_ps), ref, 2e-5, "vs torch fp32 on the same bf16 operands")
    assert_close(nchw(y_pt), torch.tanh(ref) * 0.5 + 0.5, 2e-5, "tanh + affine epilogue")
    assert_close(nchw(y_p), F.relu(ref), 6e-3, "bf16 output")


def test_small_grids_keep_the_one_tile_kernel():
    from music_style_transfer_ldm_b200 import ops
    lib = _lib()
    x = torch.randn((64, 16, 16, 32), device=DEV).bfloat16()
    wp = torch.randn((9, 64, 32), device=DEV).bfloat16()
    ops.conv(x, wp, None, ksize=3, backend=ops.BACKEND_TCGEN05)          # the UNet's enc1 at batch 64: 128 tiles
    assert lib.mstk_debug_tc_last_config()[1] >= 1
