"""-m gpu: the tcgen05 weight-gradient kernel (csrc/wgrad_tc.cu) against torch autograd (CPU fp32) on
bf16-representable operands: every (swap, taps-per-CTA, swizzle, stride, phase) combination the training
step uses, plus ragged batches and multi-split-K sizes."""
import math

import pytest
import torch

from gpu_util import assert_close, conv_ref, gen, nhwc

pytestmark = pytest.mark.gpu
DEV = "cuda"

# n, h, w, cin, cout, k, stride, transposed
CASES = [
    (4, 8, 8, 64, 128, 3, 1, False),      # M = cout, N = 64 (SW128), 3 groups of 3 taps
    (3, 16, 16, 64, 128, 3, 2, False),    # stride-2 parity views (style_encoder.enc2)
    (5, 8, 8, 128, 256, 3, 2, False),     # N = 128, two M tiles (style_encoder.enc3), ragged batch tile
    (2, 8, 8, 128, 64, 4, 2, True),       # swap: M = cin, N = cout; 4 phases x 4 taps (decoder.3)
    (2, 4, 4, 32, 128, 4, 2, True),       # N = 32 -> SWIZZLE_64B operand (decoder.0)
    (3, 4, 4, 256, 128, 3, 2, True),      # convT k3: phases with 1/2/2/4 taps (unet.dec3)
    (7, 2, 2, 512, 512, 3, 1, False),     # BNt = 256, one tap per CTA, tiny maps (unet.bottleneck)
    (37, 1, 1, 128, 128, 1, 1, False),    # Linear (time MLP)
    (9, 4, 4, 256, 512, 1, 1, False),     # attention projection as k1 conv
    (40, 16, 16, 64, 128, 3, 1, False),   # many pixel tiles -> split-K > 1
    (4, 16, 16, 32, 64, 3, 1, False),     # M side of 64 channels (upper half of the tile = out-of-range chunk), N = 32 (unet.enc1)
    (4, 16, 16, 64, 32, 3, 1, False),     # swap with a 64-channel M side (unet.dec1)
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(map(str, c)))
def test_wgrad_tc(case):
    from music_style_transfer_ldm_b200 import ops
    n, h, w, cin, cout, k, s, tr = case
    dt = torch.bfloat16
    g = gen(hash(case) % 991)
    x = torch.randn((n, cin, h, w), generator=g).to(dt).float().requires_grad_(True)
    wt = (torch.randn((cin, cout, k, k) if tr else (cout, cin, k, k), generator=g) / math.sqrt(cin * k * k)).float().requires_grad_(True)
    y = conv_ref(x, wt[:, :, 0, 0] if k == 1 else wt, None, k, s, tr)
    dy = torch.randn(y.shape, generator=g).to(dt).float()
    y.backward(dy)
    dw = torch.zeros(wt.shape, device=DEV)
    xd, dyd = nhwc(x.detach(), dt).to(DEV), nhwc(dy, dt).to(DEV)
    ops.wgrad(xd, dyd, dw, ksize=k, stride=s, transposed=tr)
    ops.wgrad(xd, dyd, dw, ksize=k, stride=s, transposed=tr)       # accumulates into dw
    assert_close(dw.cpu() / 2, wt.grad, 2e-4, "wgrad(tcgen05)")
