"""-m gpu: every backward / optimizer kernel against torch autograd (CPU, fp32) on the same operands."""
import math

import pytest
import torch
import torch.nn.functional as F

from gpu_util import assert_close, conv_ref, gen, nchw, nhwc, rel_l2

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ops():
    from music_style_transfer_ldm_b200 import ops
    return ops


# n, h, w, cin, cout, k, stride, transposed
GRAD_CASES = [(2, 8, 8, 32, 64, 3, 1, False), (2, 8, 8, 64, 32, 3, 2, False), (3, 4, 4, 64, 32, 3, 2, True), (2, 8, 8, 32, 64, 4, 2, True),
              (2, 4, 4, 64, 96, 1, 1, False), (2, 16, 16, 1, 16, 3, 2, False), (2, 8, 8, 16, 1, 4, 2, True), (1, 6, 10, 20, 12, 3, 1, False)]


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("case", GRAD_CASES, ids=lambda c: "x".join(map(str, c)))
def test_wgrad_and_dgrad(case, dt):
    ops = _ops()
    n, h, w, cin, cout, k, s, tr = case
    g = gen(hash(case) % 997)
    x = torch.randn((n, cin, h, w), generator=g).to(dt).float().requires_grad_(True)
    wt = (torch.randn((cin, cout, k, k) if tr else (cout, cin, k, k), generator=g) / math.sqrt(cin * k * k)).to(dt).float().requires_grad_(True)
    y = conv_ref(x, wt[:, :, 0, 0] if k == 1 else wt, None, k, s, tr)
    dy = torch.randn(y.shape, generator=g).to(dt).float()
    y.backward(dy)
    tol = 2e-5 if dt == torch.float32 else 1e-5      # operands are exactly representable; fp32 accumulation on both sides
    # ---- weight gradient ----
    dw = torch.zeros(wt.shape, device=DEV)
    ops.wgrad(nhwc(x.detach(), dt).to(DEV), nhwc(dy, dt).to(DEV), dw, ksize=k, stride=s, transposed=tr)
    ops.wgrad(nhwc(x.detach(), dt).to(DEV), nhwc(dy, dt).to(DEV), dw, ksize=k, stride=s, transposed=tr)   # accumulates
    assert_close(dw.cpu() / 2, wt.grad, tol * 10, "wgrad")
    # ---- data gradient through the forward kernels on a re-packed weight ----
    dyd = nhwc(dy, dt).to(DEV)
    wd = wt.detach().to(DEV)
    if tr:      # convT fwd -> conv (same k, stride 2) with the weight read as [co=cin_T][ci=cout_T]
        wp = ops.pack_weight(wd, dt)
        dx = ops.conv(dyd, wp, None, ksize=k, stride=2, out_dtype=torch.float32)
    elif s == 2:  # conv s2 fwd -> convT s2 with the weight read as [cin_T=cout][cout_T=cin]
        wp = ops.pack_weight(wd, dt, transposed=True)
        dx = ops.conv(dyd, wp, None, ksize=k, stride=2, transposed=True, out_dtype=torch.float32)
    else:
        wp = ops.pack_weight_dgrad(wd[:, :, 0, 0].contiguous() if k == 1 else wd, dt)
        dx = ops.conv(dyd, wp, None, ksize=k, stride=1, out_dtype=torch.float32)
    assert_close(nchw(dx), x.grad, tol * 10, "dgrad")


@pytest.mark.parametrize("n,h,w,dy_dt,tol", [(2, 16, 32, torch.float32, 2e-4), (3, 5, 16, torch.bfloat16, 1e-4), (1, 64, 64, torch.float32, 2e-4)])
def test_wgrad_convt4_cout1_mma(n, h, w, dy_dt, tol):
    """the decoder's last layer (ConvTranspose2d(64, 1, 4, 2, 1)): mma.sync skinny-GEMM weight gradient, bf16 x with an fp32
    (hi + lo bf16 split) or bf16 dy, against autograd; accumulation into dw checked by calling twice"""
    ops = _ops()
    g = gen(n * 7 + h)
    x = torch.randn((n, 64, h, w), generator=g).bfloat16().float()
    wt = (torch.randn((64, 1, 4, 4), generator=g) / 32).requires_grad_(True)
    dy = torch.randn((n, 1, 2 * h, 2 * w), generator=g).to(dy_dt).float()
    F.conv_transpose2d(x, wt, None, stride=2, padding=1).backward(dy)
    dw = torch.zeros(wt.shape, device=DEV)
    for _ in range(2):
        ops.wgrad(nhwc(x, torch.bfloat16).to(DEV), nhwc(dy, dy_dt).to(DEV), dw, ksize=4, stride=2, transposed=True)
    assert_close(dw.cpu() / 2, wt.grad, tol, "wgrad convT k4 cout1 (mma)")


@pytest.mark.parametrize("n,h,w,stride,x_dt,tol", [(2, 32, 64, 2, torch.float32, 2e-4), (3, 10, 32, 2, torch.bfloat16, 1e-4),
                                                   (2, 6, 16, 1, torch.float32, 2e-4), (1, 128, 128, 2, torch.float32, 2e-4)])
def test_wgrad_conv_cin1_mma(n, h, w, stride, x_dt, tol):
    """Conv2d(1, 64, 3, stride, 1) (style encoder / VAE encoder first layer): the nine-tap form of the mma.sync edge wgrad"""
    ops = _ops()
    g = gen(n * 5 + h + stride)
    x = torch.rand((n, 1, h, w), generator=g).to(x_dt).float()
    wt = (torch.randn((64, 1, 3, 3), generator=g) / 3).requires_grad_(True)
    y = F.conv2d(x, wt, None, stride=stride, padding=1)
    dy = torch.randn(y.shape, generator=g).bfloat16().float()
    y.backward(dy)
    dw = torch.zeros(wt.shape, device=DEV)
    for _ in range(2):
        ops.wgrad(nhwc(x, x_dt).to(DEV), nhwc(dy, torch.bfloat16).to(DEV), dw, ksize=3, stride=stride)
    assert_close(dw.cpu() / 2, wt.grad, tol, "wgrad conv k3 cin1 (mma)")


@pytest.mark.parametrize("c,act", [(64, 1), (32, 2), (1, 2), (24, 0), (12, 1)])
def test_act_bwd(c, act):
    ops = _ops()
    g = gen(c + act)
    rows = 777
    v = torch.randn((rows, c), generator=g, requires_grad=True)
    y = F.relu(v) if act == 1 else (torch.tanh(v) * 0.5 + 0.5 if act == 2 else v * 1.0)
    dy, dy2 = torch.randn((rows, c), generator=g), torch.randn((rows, c), generator=g)
    y.backward(dy + dy2)
    db = torch.zeros(c, device=DEV)
    got = ops.act_bwd(dy.to(DEV), y.detach().to(DEV), act=act, dy2=dy2.to(DEV), dbias=db,
                      y_scale=0.5 if act == 2 else 1.0, y_shift=0.5 if act == 2 else 0.0)
    assert_close(got, v.grad, 1e-5, "act_bwd g")
    assert_close(db, v.grad.sum(0), 1e-4, "act_bwd dbias")
    # bf16 tensors, fp32 gradient out
    got2 = ops.act_bwd(dy.bfloat16().to(DEV), y.detach().to(DEV), act=act, g_dtype=torch.float32,
                       y_scale=0.5 if act == 2 else 1.0, y_shift=0.5 if act == 2 else 0.0)
    assert rel_l2(got2, v.grad * 0 + torch.autograd.grad(
        (F.relu(v) if act == 1 else (torch.tanh(v) * 0.5 + 0.5 if act == 2 else v * 1.0)), v, dy.bfloat16().float())[0]) < 3e-5   # fp32: 1 - t*t cancels near |t| = 1


@pytest.mark.parametrize("c,relu", [(64, True), (32, False)])
def test_bn_bwd(c, relu):
    ops = _ops()
    g = gen(c)
    n, h, w = 3, 6, 5
    v = torch.randn((n, c, h, w), generator=g, requires_grad=True)
    gamma = (1 + 0.3 * torch.randn(c, generator=g)).requires_grad_(True)
    beta = (0.2 * torch.randn(c, generator=g)).requires_grad_(True)
    out = F.batch_norm(v, None, None, gamma, beta, training=True, eps=1e-5)
    out = F.relu(out) if relu else out
    dy = torch.randn(out.shape, generator=g)
    out.backward(dy)
    mean = v.detach().mean(dim=(0, 2, 3))
    invstd = torch.rsqrt(v.detach().var(dim=(0, 2, 3), unbiased=False) + 1e-5)
    dg, dbt = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    dv = ops.bn_bwd(nhwc(dy).to(DEV), nhwc(v.detach()).to(DEV), mean.to(DEV), invstd.to(DEV), gamma.detach().to(DEV), beta.detach().to(DEV),
                    act=1 if relu else 0, dgamma=dg, dbeta=dbt)
    assert_close(nchw(dv), v.grad, 2e-5, "bn dv")
    assert_close(dg, gamma.grad, 2e-5, "bn dgamma")
    assert_close(dbt, beta.grad, 2e-5, "bn dbeta")


@pytest.mark.parametrize("b,l,s,e", [(3, 16, 16, 256), (2, 4, 4, 512), (1, 48, 48, 256)])
def test_attention_bwd(b, l, s, e):
    ops = _ops()
    heads, g = 4, gen(b + l + e)
    d = e // heads
    q = torch.randn((b, l, e), generator=g, requires_grad=True)
    kv = torch.randn((b, s, 2 * e), generator=g, requires_grad=True)
    k, v = kv[:, :, :e], kv[:, :, e:]
    qh = q.view(b, l, heads, d).transpose(1, 2) / math.sqrt(d)
    kh, vh = k.reshape(b, s, heads, d).transpose(1, 2), v.reshape(b, s, heads, d).transpose(1, 2)
    o = (torch.softmax(qh @ kh.transpose(-1, -2), dim=-1) @ vh).transpose(1, 2).reshape(b, l, e)
    do = torch.randn(o.shape, generator=g)
    o.backward(do)
    kvd = kv.detach().to(DEV)
    dq, dkv = ops.attention_bwd(q.detach().to(DEV), kvd[:, :, :e], kvd[:, :, e:], do.to(DEV), heads)
    assert_close(dq, q.grad, 2e-5, "dq")
    assert_close(dkv, kv.grad, 2e-5, "dkv")


def test_time_mlp_pieces_and_small_ops():
    ops = _ops()
    g = gen(5)
    t = torch.tensor([0.0, 57.0, 199.0, 3.0, 120.0])
    w1, b1, w2, b2 = (torch.randn((128, 128), generator=g) * 0.09, torch.randn(128, generator=g) * 0.05,
                      torch.randn((128, 128), generator=g) * 0.09, torch.randn(128, generator=g) * 0.05)
    from oracle import ldm_oracle as O
    e_ref = O.sinusoidal_embedding(t, 128)
    pre_ref = F.linear(e_ref, w1, b1).requires_grad_(True)
    out_ref = F.linear(F.gelu(pre_ref), w2, b2)
    out, e, pre, hh = ops.time_embed_train(t.to(DEV), w1.to(DEV), b1.to(DEV), w2.to(DEV), b2.to(DEV))
    assert_close(e, e_ref, 1e-5, "sinusoid"); assert_close(pre, pre_ref, 1e-5, "pre"); assert_close(out, out_ref, 1e-5, "out")
    assert_close(hh, F.gelu(pre_ref.detach()), 1e-5, "gelu(pre)")
    dh = torch.randn(pre_ref.shape, generator=g)
    F.gelu(pre_ref).backward(dh)
    assert_close(ops.gelu_bwd(dh.to(DEV), pre_ref.detach().to(DEV)), pre_ref.grad, 1e-5, "gelu'")
    a, a2 = torch.randn((3, 5, 7, 24), generator=g), torch.randn((3, 5, 7, 24), generator=g)
    assert_close(ops.sum_hw(a.to(DEV), a2.to(DEV)), (a + a2).sum(dim=(1, 2)), 1e-5, "sum_hw")
    nc = torch.randn((3, 24), generator=g)
    assert_close(ops.add_bcast(a.to(DEV), a2.to(DEV), nc.to(DEV)), a + a2 + nc[:, None, None, :], 1e-6, "add_bcast")
    assert_close(ops.mse_grad(a.to(DEV), a2.to(DEV), 0.25, torch.float32), 0.25 * (a - a2), 1e-6, "mse_grad")
    ab = O.noise_schedule(200)[2]
    tt = torch.tensor([3, 150, 77])
    eh, ep, dx0 = (torch.randn((3, 4, 4, 8), generator=g) for _ in range(3))
    coef = (-torch.sqrt(1 - ab[tt]) / torch.sqrt(ab[tt])).view(3, 1, 1, 1)
    assert_close(ops.eps_grad(dx0.to(DEV), eh.to(DEV), ep.to(DEV), ab.to(DEV), tt.to(DEV), 0.5, torch.float32), 0.5 * (eh - ep) + coef * dx0, 1e-5, "eps_grad")


def test_adam_matches_torch():
    ops = _ops()
    g = gen(9)
    n = 10007
    p0, grads = torch.randn(n, generator=g), [torch.randn(n, generator=g) * (0.1 + i) for i in range(5)]
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([ref], lr=5e-4)
    p = torch.empty(n + 1, device=DEV)[:n].copy_(p0) if False else p0.clone().to(DEV)
    m, v = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    for i, gr in enumerate(grads):
        ref.grad = gr.clone(); opt.step()
        ops.adam_step(p, gr.to(DEV), m, v, lr=5e-4, beta1=0.9, beta2=0.999, eps=1e-8, step=i + 1)
    assert_close(p, ref.detach(), 1e-6, "adam params after 5 steps")


def test_pack_multi_matches_the_single_packs():
    """ops.PackPlan (one launch for every packed copy) against pack_weight / pack_weight_dgrad, all three modes, ragged tiles"""
    from music_style_transfer_ldm_b200 import ops
    g = torch.Generator().manual_seed(9)
    ws = {"conv": torch.randn((70, 33, 3, 3), generator=g).cuda(), "lin": torch.randn((130, 96), generator=g).cuda(),
          "convt": torch.randn((40, 65, 4, 4), generator=g).cuda(), "dg": torch.randn((64, 48, 3, 3), generator=g).cuda(),
          "k1": torch.randn((256, 512, 1, 1), generator=g).cuda()}
    ops.PACK_RECORDER = []
    try:
        outs = [ops.pack_weight(ws["conv"], torch.bfloat16), ops.pack_weight(ws["lin"], torch.float32),
                ops.pack_weight(ws["convt"], torch.bfloat16, transposed=True), ops.pack_weight_dgrad(ws["dg"], torch.bfloat16),
                ops.pack_weight(ws["k1"], torch.bfloat16), ops.pack_weight_dgrad(ws["lin"], torch.float32)]
        rec = ops.PACK_RECORDER
    finally:
        ops.PACK_RECORDER = None
    assert len(rec) == len(outs)
    want = [o.clone() for o in outs]
    for o in outs:
        o.zero_()
    plan = ops.PackPlan(rec)
    plan.run()
    torch.cuda.synchronize()
    for o, w_ in zip(outs, want):
        assert torch.equal(o, w_)
