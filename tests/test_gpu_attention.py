"""-m gpu: the attention cores (forward and backward) against a plain PyTorch fp32 reference, through both
kernel families: the block kernels (aligned rows, the UNet's L, S <= 16 shapes; fp32 and bf16) and the
warp-per-query kernels they fall back to (rows that are not 16-byte aligned, shapes beyond the shared-memory block)."""
import math

import pytest
import torch

from gpu_util import assert_close, gen

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ops():
    from music_style_transfer_ldm_b200 import ops
    return ops


def _ref(q, k, v, heads):
    b, l, e = q.shape
    s, d = k.shape[1], e // heads
    qh = q.view(b, l, heads, d).transpose(1, 2) / math.sqrt(d)
    kh, vh = k.reshape(b, s, heads, d).transpose(1, 2), v.reshape(b, s, heads, d).transpose(1, 2)
    return (torch.softmax(qh @ kh.transpose(-1, -2), dim=-1) @ vh).transpose(1, 2).reshape(b, l, e)


# b, l, s, e, heads
CASES = [(64, 16, 16, 256, 4),     # cross_attention2 at batch 64 (block kernels)
         (5, 4, 4, 512, 4),        # cross_attention1
         (2, 16, 4, 256, 4),       # L != S
         (3, 20, 12, 96, 4),       # d = 24: two query blocks, ragged second one
         (2, 16, 16, 40, 4),       # d = 10: rows not 16-byte aligned -> warp-per-query kernels
         (1, 128, 128, 256, 4)]    # 128x1024 spectrogram latent: beyond the block kernel's shared memory


@pytest.mark.parametrize("b,l,s,e,heads", CASES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_attention_fwd_bwd(b, l, s, e, heads, dtype):
    ops = _ops()
    g = gen(b * 7 + l + e)
    q = torch.randn((b, l, e), generator=g).to(dtype).float().requires_grad_(True)
    kv = torch.randn((b, s, 2 * e), generator=g).to(dtype).float().requires_grad_(True)
    o = _ref(q, kv[:, :, :e], kv[:, :, e:], heads)
    do = torch.randn(o.shape, generator=g).to(dtype).float()
    o.backward(do)
    tol_f, tol_b = (1e-5, 3e-5) if dtype == torch.float32 else (1e-2, 2e-2)    # bf16: output rounding of O(1) values
    qd, kvd, dod = q.detach().to(DEV, dtype), kv.detach().to(DEV, dtype), do.to(DEV, dtype)
    got = ops.attention(qd, kvd[:, :, :e], kvd[:, :, e:], heads)
    assert got.dtype == dtype
    assert_close(got, o.detach(), tol_f, "attention forward")
    dq, dkv = ops.attention_bwd(qd, kvd[:, :, :e], kvd[:, :, e:], dod, heads)
    assert_close(dq, q.grad, tol_b, "dq")
    assert_close(dkv, kv.grad, tol_b, "dkv")
