"""-m gpu: every fp32 CUDA-core kernel of libmstk against a plain PyTorch fp32 (CPU)
reference of the same op, through the C ABI (ops.* -> ctypes -> libmstk.so)."""
import math

import pytest
import torch
import torch.nn.functional as F

from gpu_util import assert_close, conv_ref, gen, nchw, nhwc, pack_ref

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _ops():
    from music_style_transfer_ldm_b200 import ops
    return ops


@pytest.fixture(autouse=True)
def _cuda_core_fp32():
    """this file pins the fp32 CUDA-core kernels; the split-operand tensor-core route of fp32 mode has its own
    tests (test_gpu_fp32_tc.py)"""
    ops = _ops()
    old, ops.FP32_TC = ops.FP32_TC, False
    yield
    ops.FP32_TC = old


CONV_CASES = [
    # n, h, w, cin, cout, k, stride, transposed
    (2, 16, 16, 32, 64, 3, 1, False),
    (2, 16, 16, 64, 128, 3, 2, False),
    (3, 8, 8, 128, 256, 3, 2, False),
    (1, 4, 4, 256, 512, 3, 2, False),
    (2, 2, 2, 512, 512, 3, 1, False),
    (2, 2, 2, 512, 256, 3, 2, True),
    (2, 8, 8, 128, 64, 3, 2, True),
    (2, 16, 16, 32, 128, 4, 2, True),
    (2, 32, 32, 64, 1, 4, 2, True),       # decoder last layer, Cout = 1
    (2, 64, 64, 1, 64, 3, 2, False),      # first layers, Cin = 1
    (2, 16, 24, 1, 64, 3, 1, False),      # VGGish first layer
    (2, 4, 4, 256, 256, 1, 1, False),     # linear (attention projection)
    (1, 16, 48, 32, 64, 3, 1, False),     # non power-of-two width (config 5 latent)
    (3, 6, 10, 20, 24, 3, 2, False),      # ragged: nothing divides the tiles
]


@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: "x".join(map(str, c)))
def test_conv_simt_fp32(case):
    ops = _ops()
    n, h, w, cin, cout, k, s, tr = case
    g = gen(hash(case) % 1000)
    x = torch.randn((n, cin, h, w), generator=g)
    wt = torch.randn((cin, cout, k, k) if tr else (cout, cin, k, k), generator=g) / math.sqrt(cin * k * k)
    b = torch.randn((cout,), generator=g)
    ref = conv_ref(x, wt[:, :, 0, 0] if k == 1 else wt, b, k, s, tr)
    wp = ops.pack_weight(wt.to(DEV), torch.float32, transposed=tr)
    assert torch.equal(wp.cpu(), pack_ref(wt, tr))
    y = ops.conv(nhwc(x).to(DEV), wp, b.to(DEV), ksize=k, stride=s, transposed=tr, backend=ops.BACKEND_SIMT)
    assert_close(nchw(y), ref, 2e-6, "conv")


def test_conv_epilogue_fp32():
    """bias -> stats -> relu -> affine -> +add_nc -> +add, and the tanh/(y+1)/2 variant"""
    ops = _ops()
    g = gen(5)
    n, h, w, cin, cout = 3, 8, 8, 16, 40
    x = torch.randn((n, cin, h, w), generator=g)
    wt = torch.randn((cout, cin, 3, 3), generator=g) / 12
    b = torch.randn((cout,), generator=g)
    add = torch.randn((n, cout, h, w), generator=g)
    add_nc = torch.randn((n, cout), generator=g)
    pre = F.conv2d(x, wt, b, padding=1)
    ref = F.relu(pre) * 0.5 + 0.25 + add_nc[:, :, None, None] + add
    stats = torch.zeros(2 * cout, dtype=torch.float64, device=DEV)
    wp = ops.pack_weight(wt.to(DEV), torch.float32)
    y = ops.conv(nhwc(x).to(DEV), wp, b.to(DEV), ksize=3, act=ops.ACT_RELU, add=nhwc(add).to(DEV), add_nc=add_nc.to(DEV),
                 stats=stats, out_scale=0.5, out_shift=0.25)
    assert_close(nchw(y), ref, 2e-6, "epilogue")
    assert_close(stats[:cout], pre.double().sum(dim=(0, 2, 3)), 1e-6, "stats sum")
    assert_close(stats[cout:], (pre.double() ** 2).sum(dim=(0, 2, 3)), 1e-6, "stats sumsq")
    y2 = ops.conv(nhwc(x).to(DEV), wp, b.to(DEV), ksize=3, act=ops.ACT_TANH, out_scale=0.5, out_shift=0.5)
    assert_close(nchw(y2), (torch.tanh(pre) + 1) / 2, 2e-6, "tanh")
    # bf16 output of an fp32 contraction
    y3 = ops.conv(nhwc(x).to(DEV), wp, b.to(DEV), ksize=3, out_dtype=torch.bfloat16)
    assert y3.dtype == torch.bfloat16
    assert_close(nchw(y3), pre, 5e-3, "bf16 out")


def test_conv_ddim_epilogue_fp32():
    ops = _ops()
    from oracle import ldm_oracle as O
    g = gen(6)
    n, h, w, cin, cout = 2, 16, 16, 64, 32
    x = torch.randn((n, cin, h, w), generator=g)
    wt = torch.randn((cout, cin, 3, 3), generator=g) / 24
    b = torch.randn((cout,), generator=g) * 0.1
    lat = torch.randn((n, cout, h, w), generator=g)
    sd = {"noise_scheduler.alpha_bar_t": O.noise_schedule(200)[2]}
    eps = F.conv2d(x, wt, b, padding=1)
    xn_ref, x0_ref = O.ddim_update(sd, lat, eps, 120, 95, 0.3)
    ab = sd["noise_scheduler.alpha_bar_t"]
    sab, s1m = torch.sqrt(ab).tolist(), torch.sqrt(1 - ab).tolist()
    xc = nhwc(lat).to(DEV)
    x0 = torch.empty_like(xc)
    xlp = torch.empty(xc.shape, dtype=torch.bfloat16, device=DEV)
    d = ops.Ddim(x_cur=xc, x_next=xc, x_next_lp=xlp, x0_pred=x0, sab_t=sab[120], s1m_t=s1m[120], sab_n=sab[95], s1m_n=s1m[95], eta=0.3)
    y = ops.conv(nhwc(x).to(DEV), ops.pack_weight(wt.to(DEV), torch.float32), b.to(DEV), ksize=3, ddim=d)
    assert_close(nchw(y), eps, 2e-6, "eps")
    assert_close(nchw(xc), xn_ref, 5e-6, "x_next (in place)")
    assert_close(nchw(x0), x0_ref, 5e-6, "x0_pred")
    assert_close(nchw(xlp), xn_ref, 5e-3, "bf16 copy")
    # eps-less variant + stand-alone kernel
    xc2 = nhwc(lat).to(DEV)
    d2 = ops.Ddim(x_cur=xc2, x_next=xc2, sab_t=sab[120], s1m_t=s1m[120], sab_n=sab[95], s1m_n=s1m[95], eta=0.3)
    assert ops.conv(nhwc(x).to(DEV), ops.pack_weight(wt.to(DEV), torch.float32), b.to(DEV), ksize=3, ddim=d2, want_y=False) is None
    assert_close(nchw(xc2), xn_ref, 5e-6, "x_next (no eps)")
    xc3 = nhwc(lat).to(DEV)
    ops.ddim_step(ops.Ddim(x_cur=xc3, x_next=xc3, sab_t=sab[120], s1m_t=s1m[120], sab_n=sab[95], s1m_n=s1m[95], eta=0.3),
                  nhwc(eps).to(DEV))
    assert_close(nchw(xc3), xn_ref, 5e-6, "ddim_step")


def test_layout_and_cast():
    ops = _ops()
    x = torch.randn((3, 37, 5, 9), generator=gen(1))
    a = ops.nchw_to_nhwc(x.to(DEV), torch.float32)
    assert torch.equal(a.cpu(), nhwc(x))
    assert torch.equal(ops.nhwc_to_nchw(a).cpu(), x)
    ab = ops.nchw_to_nhwc(x.to(DEV), torch.bfloat16)
    assert torch.equal(ab.cpu(), nhwc(x).bfloat16())
    assert torch.equal(ops.cast(a, torch.bfloat16).cpu(), nhwc(x).bfloat16())
    assert torch.equal(ops.nhwc_to_nchw(ab).cpu(), x.bfloat16().float())


@pytest.mark.parametrize("c,act", [(64, 1), (32, 0), (12, 1)])
def test_bn_apply(c, act):
    ops = _ops()
    g = gen(c)
    x = torch.randn((4, c, 6, 10), generator=g) * 2 + 0.5
    gamma, beta = torch.rand(c, generator=g) + 0.5, torch.randn(c, generator=g)
    rm, rv = torch.randn(c, generator=g), torch.rand(c, generator=g) + 0.5
    bn = torch.nn.BatchNorm2d(c)
    bn.load_state_dict({"weight": gamma, "bias": beta, "running_mean": rm, "running_var": rv, "num_batches_tracked": torch.tensor(0)})
    bn.train()
    ref = bn(x)
    if act:
        ref = F.relu(ref)
    xd = nhwc(x).to(DEV)
    stats = torch.cat((x.double().sum(dim=(0, 2, 3)), (x.double() ** 2).sum(dim=(0, 2, 3)))).to(DEV)
    rmd, rvd = rm.clone().to(DEV), rv.clone().to(DEV)
    sm, si = torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    y = ops.bn_apply(xd, stats, gamma.to(DEV), beta.to(DEV), count=4 * 6 * 10, eps=1e-5, momentum=0.1, act=act,
                     out_dtype=torch.float32, running_mean=rmd, running_var=rvd, save_mean=sm, save_invstd=si)
    assert_close(nchw(y), ref, 5e-6, "bn out")
    assert_close(rmd, bn.running_mean, 1e-6, "running_mean")
    assert_close(rvd, bn.running_var, 1e-6, "running_var")
    assert_close(sm, x.mean(dim=(0, 2, 3)), 1e-5, "save_mean")


def test_time_embed():
    ops = _ops()
    from oracle import ldm_oracle as O
    sd = O.make_state_dict(1234)
    for t in (torch.tensor([0, 57, 199, 120]), torch.tensor([37.0, 0.5])):
        ref = O.time_mlp(sd, t)
        got = ops.time_embed(t.float().to(DEV), sd["unet.time_mlp.1.weight"].to(DEV), sd["unet.time_mlp.1.bias"].to(DEV),
                             sd["unet.time_mlp.3.weight"].to(DEV), sd["unet.time_mlp.3.bias"].to(DEV))
        assert_close(got, ref, 3e-5, "time_embed")


@pytest.mark.parametrize("b,l,s,e", [(3, 16, 16, 256), (2, 4, 4, 512), (1, 128, 128, 256), (2, 48, 48, 256)])
def test_attention_core(b, l, s, e):
    ops = _ops()
    g = gen(l + e)
    heads, d = 4, e // 4
    q = torch.randn((b, l, e), generator=g)
    kv = torch.randn((b, s, 2 * e), generator=g)
    k, v = kv[:, :, :e], kv[:, :, e:]
    qh = q.view(b, l, heads, d).transpose(1, 2)
    kh = k.reshape(b, s, heads, d).transpose(1, 2)
    vh = v.reshape(b, s, heads, d).transpose(1, 2)
    ref = (torch.softmax(qh @ kh.transpose(-1, -2) / math.sqrt(d), -1) @ vh).transpose(1, 2).reshape(b, l, e)
    kvd = kv.to(DEV)
    got = ops.attention(q.to(DEV), kvd[:, :, :e], kvd[:, :, e:], heads)
    assert_close(got, ref, 1e-5, "attention")


def test_scheduler_kernels():
    ops = _ops()
    from oracle import ldm_oracle as O
    sd = {"noise_scheduler.alpha_bar_t": O.noise_schedule(200)[2]}
    g = gen(3)
    x0, eps = torch.randn((3, 32, 16, 16), generator=g), torch.randn((3, 32, 16, 16), generator=g)
    t = torch.tensor([0, 57, 199])
    ab = sd["noise_scheduler.alpha_bar_t"].to(DEV)
    z, zlp = ops.q_sample(x0.to(DEV), eps.to(DEV), ab, t.to(DEV), lp=True)
    assert_close(z, O.q_sample(sd, x0, t, eps), 1e-6, "q_sample")
    assert_close(zlp, O.q_sample(sd, x0, t, eps), 5e-3, "q_sample bf16")
    p, _ = ops.predict_x0(z, eps.to(DEV), ab, t.to(DEV))
    assert_close(p, O.predict_x0(sd, z.cpu(), t, eps), 1e-6, "predict_x0")


def test_loss_reductions():
    ops = _ops()
    from oracle import ldm_oracle as O
    g = gen(4)
    a, b = torch.randn((5, 32, 16, 16), generator=g), torch.randn((5, 32, 16, 16), generator=g)
    for _ in range(2):   # twice: the scratch must come back zeroed
        assert_close(ops.mse(a.to(DEV), b.to(DEV)), O.diffusion_loss(a, b).reshape(1), 1e-6, "mse")
        assert_close(ops.kl(a.to(DEV)), O.kl_regularization_loss(a).reshape(1), 1e-5, "kl")
    big = torch.rand((2, 1, 128, 1024), generator=g)
    assert_close(ops.mse(big.to(DEV), (big * 0.5).to(DEV)), O.diffusion_loss(big, big * 0.5).reshape(1), 1e-6, "mse big")
    f, h = torch.relu(torch.randn((3, 40, 9, 7), generator=g)), torch.relu(torch.randn((3, 40, 9, 7), generator=g)) * 3
    fn = f / (f.std(dim=[1, 2, 3], keepdim=True) + 1e-8)
    hn = h / (h.std(dim=[1, 2, 3], keepdim=True) + 1e-8)
    ref = ((fn - hn) ** 2).mean().reshape(1)
    for _ in range(2):
        out = torch.zeros(1, device=DEV)
        ops.feature_std_mse(nhwc(f).to(DEV), nhwc(h).to(DEV), 0.5, out)
        ops.feature_std_mse(nhwc(f).to(DEV), nhwc(h).to(DEV), 0.5, out)
        assert_close(out, ref, 1e-5, "feature_std_mse")


def test_maxpool():
    ops = _ops()
    x = torch.randn((2, 24, 8, 12), generator=gen(8))
    assert torch.equal(nchw(ops.maxpool2x2(nhwc(x).to(DEV))).cpu(), F.max_pool2d(x, 2, 2))


def test_errors_are_loud():
    ops = _ops()
    with pytest.raises(ops.MstkError):
        ops.conv(torch.zeros((1, 4, 4, 8)), torch.zeros((9, 8, 8)), None, ksize=3)        # CPU tensor
    x = torch.zeros((1, 4, 4, 8), device=DEV)
    with pytest.raises(ops.MstkError):
        ops.conv(x, torch.zeros((9, 8, 4), device=DEV), None, ksize=3)                      # cin mismatch
    with pytest.raises(ops.MstkError):
        ops.conv(x, torch.zeros((9, 8, 8), device=DEV), None, ksize=3, backend=ops.BACKEND_TCGEN05)   # fp32 on tcgen05
