"""-m gpu: attention cores (forward / backward, csrc/elementwise.cu + backward.cu) and the composed cross-attention layer
(csrc/xattn.cu) on seeded random token counts, head sizes and batch sizes, against plain PyTorch fp32 / the oracle's
nn.MultiheadAttention restatement (models/model.py:126-160).  The block kernels serve aligned rows and small S; everything else
falls to the warp-per-query kernels -- the draw crosses that boundary in both directions (head size not a multiple of 8, S up to
the backward kernel's limit of 128, query tiles that are not full, style maps with a different token count than the features)."""
import random

import pytest
import torch

from gpu_util import assert_close, assert_close_mode, gen
from test_gpu_attention import _ref
from test_gpu_xattn import _attn_sd, _module

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _cases(seed, count):
    rng = random.Random(seed)
    out = []
    for _ in range(count):
        heads = rng.choice([1, 2, 4, 8])
        d = rng.choice([4, 8, 10, 16, 24, 32, 64, 100])
        out.append((rng.randint(1, 5), rng.randint(1, 70), rng.randint(1, 128), heads * d, heads, rng.randrange(1 << 20)))
    return out


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16], ids=["fp32", "bf16"])
@pytest.mark.parametrize("b,l,s,e,heads,seed", _cases(5, 16), ids=lambda v: str(v))
def test_attention_random_shapes(b, l, s, e, heads, seed, dtype):
    from music_style_transfer_ldm_b200 import ops
    g = gen(seed)
    q = torch.randn((b, l, e), generator=g).to(dtype).float().requires_grad_(True)
    kv = torch.randn((b, s, 2 * e), generator=g).to(dtype).float().requires_grad_(True)
    o = _ref(q, kv[:, :, :e], kv[:, :, e:], heads)
    do = torch.randn(o.shape, generator=g).to(dtype).float()
    o.backward(do)
    tol_f, tol_b = (1e-5, 5e-5) if dtype == torch.float32 else (1e-2, 2e-2)
    qd, kvd, dod = q.detach().to(DEV, dtype), kv.detach().to(DEV, dtype), do.to(DEV, dtype)
    got = ops.attention(qd, kvd[:, :, :e], kvd[:, :, e:], heads)
    assert_close(got, o.detach(), tol_f, f"attention forward b{b} l{l} s{s} e{e} h{heads}")
    dq, dkv = ops.attention_bwd(qd, kvd[:, :, :e], kvd[:, :, e:], dod, heads)
    assert_close(dq, q.grad, tol_b, f"dq b{b} l{l} s{s} e{e} h{heads}")
    assert_close(dkv, kv.grad, tol_b, f"dkv b{b} l{l} s{s} e{e} h{heads}")


def _xcases(seed, count):
    rng = random.Random(seed)
    out = []
    for _ in range(count):
        h, w = rng.randint(1, 9), rng.randint(1, 14)
        divs = [d for d in range(1, h * w + 1) if (h * w) % d == 0]
        sh = rng.choice(divs)
        out.append((rng.randint(1, 6), rng.choice([256, 512]), (h, w), (sh, h * w // sh)))
    return out


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("b,e,hw,shw", _xcases(6, 10), ids=lambda v: str(v).replace(" ", ""))
def test_cross_attention_layer_random_token_counts(b, e, hw, shw, precision):
    """the CrossAttention module (NCHW in / out) on random token counts; the style map has the feature map's token count in
    another factorisation -- the reference reshapes it with the features' h*w (model.py:149-150), so that passes there too"""
    from oracle import ldm_oracle as O
    from music_style_transfer_ldm_b200.ops import MstkError
    sd = _attn_sd(e, 3 + e)
    g = gen(b + e + hw[0] * 31 + shw[1])
    feat = torch.relu(torch.randn((b, e, hw[0], hw[1]), generator=g))
    style = torch.relu(torch.randn((b, e, shw[0], shw[1]), generator=g))
    ref = O.cross_attention(sd, "ca", feat, style)
    ca = _module(e, sd, precision)
    got = ca(feat.to(DEV), style.to(DEV))
    assert got.shape == feat.shape
    assert_close_mode(got, ref, precision, f"cross-attention e{e} b{b} tokens {hw[0] * hw[1]} as {shw}")
    with pytest.raises(MstkError, match="tokens"):          # the reference's reshape fails for any other token count
        ca(feat.to(DEV), torch.zeros((b, e, hw[0] + 1, hw[1]), device=DEV))
