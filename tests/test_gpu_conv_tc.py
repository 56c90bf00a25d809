"""-m gpu: the tcgen05/TMEM/TMA implicit-GEMM kernel against (a) the CUDA-core kernel on the
same bf16 operands and (b) a plain PyTorch fp32 reference on the bf16-rounded operands."""
import math

import pytest
import torch
import torch.nn.functional as F

from gpu_util import assert_close, conv_ref, gen, nchw, nhwc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ops():
    from music_style_transfer_ldm_b200 import ops
    return ops


TC_CASES = [
    # n, h, w, cin, cout, k, stride, transposed          (UNet / VAE / style / VGGish layer shapes)
    (2, 16, 16, 64, 64, 3, 1, False),      # simplest: KC=64, one tap list, no split
    (2, 16, 16, 32, 64, 3, 1, False),      # enc1: KC=32 (SWIZZLE_64B)
    (4, 16, 16, 64, 128, 3, 2, False),     # enc2: stride 2 parity views
    (3, 8, 8, 128, 256, 3, 2, False),      # enc3 (batch not filling the tile)
    (2, 4, 4, 256, 512, 3, 2, False),      # enc4: split-K
    (2, 2, 2, 512, 512, 3, 1, False),      # bottleneck: split-K, tiny grid
    (2, 2, 2, 512, 256, 3, 2, True),       # dec4: conv-transpose phases
    (2, 4, 4, 256, 128, 3, 2, True),       # dec3
    (2, 8, 8, 128, 64, 3, 2, True),        # dec2
    (2, 16, 16, 64, 32, 3, 1, False),      # dec1: N = 32
    (2, 16, 16, 32, 128, 4, 2, True),      # decoder convT k4
    (2, 32, 32, 128, 64, 4, 2, True),
    (2, 64, 64, 64, 128, 3, 2, False),     # VAE encoder conv2
    (2, 32, 32, 128, 32, 3, 2, False),     # VAE encoder conv3
    (2, 4, 4, 256, 256, 1, 1, False),      # attention projection
    (2, 2, 2, 512, 1024, 1, 1, False),     # fused K/V projection
    (1, 16, 48, 64, 64, 3, 1, False),      # width 48 -> Wt = 16
    (1, 64, 64, 64, 128, 3, 1, False),     # VGGish-like, many tiles
    (65, 2, 2, 512, 512, 3, 1, False),     # batch spills into a third tile
]


@pytest.mark.parametrize("case", TC_CASES, ids=lambda c: "x".join(map(str, c)))
def test_conv_tcgen05(case):
    ops = _ops()
    n, h, w, cin, cout, k, s, tr = case
    g = gen(hash(case) % 1000)
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cin, cout, k, k) if tr else (cout, cin, k, k), generator=g) / math.sqrt(cin * k * k)).bfloat16()
    b = torch.randn((cout,), generator=g)
    ref = conv_ref(x.float(), wt.float()[:, :, 0, 0] if k == 1 else wt.float(), b, k, s, tr)
    xd = nhwc(x, torch.bfloat16).to(DEV)
    wp = ops.pack_weight(wt.float().to(DEV), torch.bfloat16, transposed=tr)
    y_tc = ops.conv(xd, wp, b.to(DEV), ksize=k, stride=s, transposed=tr, out_dtype=torch.float32, backend=ops.BACKEND_TCGEN05)
    y_simt = ops.conv(xd, wp, b.to(DEV), ksize=k, stride=s, transposed=tr, out_dtype=torch.float32, backend=ops.BACKEND_SIMT)
    torch.cuda.synchronize()
    assert_close(nchw(y_simt), ref, 1e-5, "simt-bf16 vs torch")
    assert_close(nchw(y_tc), ref, 1e-5, "tcgen05 vs torch")
    # run again: the split-K workspace must have been left zeroed
    y_tc2 = ops.conv(xd, wp, b.to(DEV), ksize=k, stride=s, transposed=tr, out_dtype=torch.float32, backend=ops.BACKEND_TCGEN05)
    assert_close(nchw(y_tc2), ref, 1e-5, "tcgen05 second launch")


def test_conv_tcgen05_epilogue():
    ops = _ops()
    g = gen(11)
    n, h, w, cin, cout = 4, 8, 8, 64, 128
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cout, cin, 3, 3), generator=g) / 24).bfloat16()
    b = torch.randn((cout,), generator=g)
    add = torch.randn((n, cout, h, w), generator=g).bfloat16()
    add_nc = torch.randn((n, cout), generator=g)
    pre = F.conv2d(x.float(), wt.float(), b, padding=1)
    ref = F.relu(pre) + add_nc[:, :, None, None] + add.float()
    stats = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    wp = ops.pack_weight(wt.float().to(DEV), torch.bfloat16)
    y = ops.conv(nhwc(x, torch.bfloat16).to(DEV), wp, b.to(DEV), ksize=3, act=ops.ACT_RELU, add=nhwc(add, torch.bfloat16).to(DEV),
                 add_nc=add_nc.to(DEV), stats=stats, backend=ops.BACKEND_TCGEN05)
    assert y.dtype == torch.bfloat16
    assert_close(nchw(y), ref, 8e-3, "epilogue (bf16 out)")
    assert_close(stats[:cout], pre.double().sum(dim=(0, 2, 3)), 1e-4, "stats sum")
    assert_close(stats[cout:], (pre.double() ** 2).sum(dim=(0, 2, 3)), 1e-4, "stats sumsq")


def test_conv_tcgen05_ddim():
    ops = _ops()
    from oracle import ldm_oracle as O
    g = gen(12)
    n, h, w, cin, cout = 3, 16, 16, 64, 32
    x = torch.randn((n, cin, h, w), generator=g).bfloat16()
    wt = (torch.randn((cout, cin, 3, 3), generator=g) / 24).bfloat16()
    b = torch.randn((cout,), generator=g) * 0.1
    lat = torch.randn((n, cout, h, w), generator=g)
    sd = {"noise_scheduler.alpha_bar_t": O.noise_schedule(200)[2]}
    eps = F.conv2d(x.float(), wt.float(), b, padding=1)
    xn_ref, _ = O.ddim_update(sd, lat, eps, 49, 48, 0.0)
    ab = sd["noise_scheduler.alpha_bar_t"]
    sab, s1m = torch.sqrt(ab).tolist(), torch.sqrt(1 - ab).tolist()
    xc = nhwc(lat).to(DEV)
    xlp = torch.empty(xc.shape, dtype=torch.bfloat16, device=DEV)
    d = ops.Ddim(x_cur=xc, x_next=xc, x_next_lp=xlp, sab_t=sab[49], s1m_t=s1m[49], sab_n=sab[48], s1m_n=s1m[48], eta=0.0)
    wp = ops.pack_weight(wt.float().to(DEV), torch.bfloat16)
    y = ops.conv(nhwc(x, torch.bfloat16).to(DEV), wp, b.to(DEV), ksize=3, ddim=d, out_dtype=torch.float32, backend=ops.BACKEND_TCGEN05)
    assert_close(nchw(y), eps, 1e-5, "eps")
    assert_close(nchw(xc), xn_ref, 1e-5, "x_next")
    assert_close(nchw(xlp), xn_ref, 5e-3, "x_next bf16")


def test_conv_tcgen05_cluster_splitk():
    """opt-in MSTK_TC_CLUSTER=1 path (split-K partners reduce through distributed shared memory): the environment is
    read once per process, so the split-K cases run in a child interpreter"""
    import os, subprocess, sys
    env = dict(os.environ, MSTK_TC_CLUSTER="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_conv_tc.py"), "-q", "-x",
                        "-k", "2x4x4x256x512 or 2x2x2x512x512 or 2x2x2x512x256 or 65x2x2 or epilogue or ddim"],
                       env=env, cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
