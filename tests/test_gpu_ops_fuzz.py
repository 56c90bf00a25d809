"""-m gpu: the element-wise, layout, reduction and normalisation kernels (csrc/elementwise.cu, backward.cu) on seeded random RAGGED
sizes and on contiguous views that start 1-3 elements into an allocation (so their base address is not 16-byte aligned), against
PyTorch on the CPU.  These kernels are "128-bit vectorised where alignment allows": the layer-shape tests only ever hand them
512-byte-aligned allocations with channel counts that are multiples of 8, so the scalar tails and the alignment fall-backs never
ran.  A kernel may also REFUSE a misaligned pointer (MstkError naming the alignment) -- what it may not do is fault."""
import random

import pytest
import torch
import torch.nn.functional as F

from gpu_util import gen, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"
BF, F32 = torch.bfloat16, torch.float32


def _ops():
    from music_style_transfer_ldm_b200 import ops
    return ops


def dev(t: torch.Tensor, off: int = 0) -> torch.Tensor:
    """a contiguous CUDA copy of t whose first element sits `off` elements into its allocation"""
    buf = torch.empty(t.numel() + off, dtype=t.dtype, device=DEV)
    v = buf[off:off + t.numel()].view(t.shape)
    v.copy_(t)
    return v


def _guard(fn, off, what):
    """runs fn(); with a misaligned view a loud refusal naming the alignment is accepted (returns None)"""
    from music_style_transfer_ldm_b200.ops import MstkError
    try:
        return fn()
    except MstkError as e:
        assert off != 0 and "align" in str(e).lower(), f"{what}: {e}"
        return None


def _rng(seed):
    return random.Random(seed), gen(seed)


SEEDS = list(range(12))


@pytest.mark.parametrize("seed", SEEDS)
def test_cast_and_layout(seed):
    ops = _ops()
    r, g = _rng(seed)
    off = r.choice([0, 0, 1, 2, 3])
    n = r.choice([1, 2, 7, 31, 257, 1000, 4099, r.randint(1, 70000)])
    x = torch.randn(n, generator=g)
    got = _guard(lambda: ops.cast(dev(x, off), BF), off, "cast f32->bf16")
    if got is not None:
        assert torch.equal(got.cpu(), x.bfloat16())
    got = _guard(lambda: ops.cast(dev(x.bfloat16(), off), F32), off, "cast bf16->f32")
    if got is not None:
        assert torch.equal(got.cpu(), x.bfloat16().float())
    nn_, c, h, w = r.randint(1, 4), r.choice([1, 2, 3, 8, 32, 33, 65]), r.randint(1, 19), r.randint(1, 23)
    y = torch.randn((nn_, c, h, w), generator=g)
    for dt in (F32, BF):
        a = _guard(lambda: ops.nchw_to_nhwc(dev(y, off), dt), off, "nchw_to_nhwc")
        if a is not None:
            assert torch.equal(a.cpu(), y.permute(0, 2, 3, 1).contiguous().to(dt))
        b = _guard(lambda: ops.nhwc_to_nchw(dev(y.permute(0, 2, 3, 1).contiguous().to(dt), off)), off, "nhwc_to_nchw")
        if b is not None:
            assert torch.equal(b.cpu(), y.to(dt).float())


@pytest.mark.parametrize("seed", SEEDS)
def test_loss_reductions_and_their_gradients(seed):
    ops = _ops()
    r, g = _rng(100 + seed)
    off = r.choice([0, 0, 1, 2, 3])
    n = r.choice([1, 3, 8, 255, 1025, 65537, r.randint(1, 300000)])
    a, b = torch.randn(n, generator=g), torch.randn(n, generator=g)
    for dt, tol in ((F32, 2e-6), (BF, 2e-6)):
        aa, bb = a.to(dt), b.to(dt)
        want = ((aa.double() - bb.double()) ** 2).mean().reshape(1)
        for _ in range(2):                                            # twice: the reduction scratch must come back zeroed
            got = _guard(lambda: ops.mse(dev(aa, off), dev(bb, off)), off, "mse")
            if got is not None:
                assert rel_err(got, want) <= tol, f"mse n={n} off={off} {dt}"
    z = a + 0.1 * torch.sign(a)                                       # keep log(z^2) finite
    from oracle import ldm_oracle as O
    want = O.kl_regularization_loss(z.double()).reshape(1)
    got = _guard(lambda: ops.kl(dev(z, off)), off, "kl")
    if got is not None:
        assert rel_err(got, want) <= 1e-5, f"kl n={n} off={off}"
    got = _guard(lambda: ops.mse_grad(dev(a, off), dev(b, off), 0.37, F32), off, "mse_grad")
    if got is not None:
        assert rel_err(got, 0.37 * (a - b)) <= 1e-6
    dz = torch.randn(n, generator=g)
    got = _guard(lambda: ops.kl_grad(dev(z, off), dev(dz, off), 0.01), off, "kl_grad")
    if got is not None:
        assert rel_err(got, dz + 0.01 * (z - z / (z * z + 1e-8))) <= 1e-5


@pytest.mark.parametrize("seed", SEEDS)
def test_row_channel_kernels(seed):
    """act_bwd (+ bias sums), add_bcast, sum_hw, maxpool2x2 on [n, h, w, c] tensors with arbitrary c"""
    ops = _ops()
    r, g = _rng(200 + seed)
    off = r.choice([0, 0, 1, 2, 3])
    n, h, w = r.randint(1, 4), r.randint(1, 13), r.randint(1, 17)
    c = r.choice([1, 2, 3, 5, 8, 12, 24, 32, 33, 64, 100, 128, 257])
    act = r.choice([0, 1, 2])
    v = torch.randn((n, h, w, c), generator=g, requires_grad=True)
    y = F.relu(v) if act == 1 else (torch.tanh(v) * 0.5 + 0.5 if act == 2 else v * 1.0)
    dy, dy2 = torch.randn(v.shape, generator=g), torch.randn(v.shape, generator=g)
    y.backward(dy + dy2)
    db = torch.zeros(c, device=DEV)
    kw = dict(y_scale=0.5 if act == 2 else 1.0, y_shift=0.5 if act == 2 else 0.0)
    got = _guard(lambda: ops.act_bwd(dev(dy, off), dev(y.detach(), off), act=act, dy2=dev(dy2, off), dbias=db, **kw), off, "act_bwd")
    if got is not None:
        assert rel_err(got, v.grad) <= 2e-5, f"act_bwd c={c} act={act} off={off}"
        assert rel_err(db, v.grad.sum(dim=(0, 1, 2))) <= 1e-4, f"act_bwd dbias c={c} act={act} off={off}"
    db2 = torch.zeros(c, device=DEV)
    if _guard(lambda: ops.act_bwd(dev(dy.bfloat16(), off), None, act=0, dbias=db2, want_g=False) or True, off, "act_bwd bias only"):
        assert rel_err(db2, dy.bfloat16().float().sum(dim=(0, 1, 2))) <= 1e-4, f"bias-only sums c={c} off={off}"
    a, a2, nc = torch.randn(v.shape, generator=g), torch.randn(v.shape, generator=g), torch.randn((n, c), generator=g)
    for dt, tol in ((F32, 1e-6), (BF, 8e-3)):
        got = _guard(lambda: ops.add_bcast(dev(a.to(dt), off), dev(a2.to(dt), off), dev(nc, off)), off, "add_bcast")
        if got is not None:
            assert rel_err(got, a.to(dt).float() + a2.to(dt).float() + nc[:, None, None, :]) <= tol, f"add_bcast c={c} off={off} {dt}"
        got = _guard(lambda: ops.sum_hw(dev(a.to(dt), off), dev(a2.to(dt), off)), off, "sum_hw")
        if got is not None:
            assert rel_err(got, (a.to(dt).float() + a2.to(dt).float()).sum(dim=(1, 2))) <= 2e-5, f"sum_hw c={c} off={off} {dt}"
        if h >= 2 and w >= 2:
            got = _guard(lambda: ops.maxpool2x2(dev(a.to(dt), off)), off, "maxpool2x2")
            if got is not None:
                want = F.max_pool2d(a.to(dt).float().permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)
                assert torch.equal(got.float().cpu(), want), f"maxpool c={c} {h}x{w} off={off} {dt}"


@pytest.mark.parametrize("seed", SEEDS)
def test_batchnorm_forward_and_backward(seed):
    ops = _ops()
    r, g = _rng(300 + seed)
    off = r.choice([0, 0, 1, 2, 3])
    n, h, w = r.randint(1, 5), r.randint(1, 11), r.randint(2, 13)
    c = r.choice([1, 3, 8, 12, 32, 33, 64, 100, 128])
    relu = r.random() < 0.5
    x = (torch.randn((n, c, h, w), generator=g) * 2 + 0.5).requires_grad_(True)
    gamma = (torch.rand(c, generator=g) + 0.5).requires_grad_(True)
    beta = torch.randn(c, generator=g).requires_grad_(True)
    rm, rv = torch.randn(c, generator=g), torch.rand(c, generator=g) + 0.5
    rm_ref, rv_ref = rm.clone(), rv.clone()
    ref = F.batch_norm(x, rm_ref, rv_ref, gamma, beta, training=True, momentum=0.1, eps=1e-5)
    ref = F.relu(ref) if relu else ref
    dy = torch.randn(ref.shape, generator=g)
    ref.backward(dy)
    xl = x.detach().permute(0, 2, 3, 1).contiguous()
    stats = torch.cat((x.detach().double().sum(dim=(0, 2, 3)), (x.detach().double() ** 2).sum(dim=(0, 2, 3)))).to(DEV)
    rmd, rvd, sm, si = rm.clone().to(DEV), rv.clone().to(DEV), torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    yv = _guard(lambda: ops.bn_apply(dev(xl, off), stats, gamma.detach().to(DEV), beta.detach().to(DEV), count=n * h * w, eps=1e-5,
                                     momentum=0.1, act=1 if relu else 0, out_dtype=F32, running_mean=rmd, running_var=rvd,
                                     save_mean=sm, save_invstd=si), off, "bn_apply")
    if yv is not None:
        assert rel_err(yv.permute(0, 3, 1, 2), ref.detach()) <= 1e-5, f"bn_apply c={c} off={off}"
        if n * h * w > 1:
            assert rel_err(rmd, rm_ref) <= 1e-5 and rel_err(rvd, rv_ref) <= 1e-5, f"running statistics c={c}"
    mean = x.detach().mean(dim=(0, 2, 3))
    invstd = torch.rsqrt(x.detach().var(dim=(0, 2, 3), unbiased=False) + 1e-5)
    dg, dbt = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    dv = None if c % 8 else _guard(lambda: ops.bn_bwd(dev(dy.permute(0, 2, 3, 1).contiguous(), off), dev(xl, off), mean.to(DEV), invstd.to(DEV),
                                   gamma.detach().to(DEV), beta.detach().to(DEV), act=1 if relu else 0, dgamma=dg, dbeta=dbt), off, "bn_bwd")
    if dv is not None:
        assert rel_err(dv.permute(0, 3, 1, 2), x.grad) <= 5e-5, f"bn_bwd dv c={c} off={off} rows={n * h * w}"
        assert rel_err(dg, gamma.grad) <= 5e-5 and rel_err(dbt, beta.grad) <= 5e-5, f"bn_bwd dgamma / dbeta c={c} off={off}"


@pytest.mark.parametrize("seed", SEEDS)
def test_scheduler_kernels_and_adam(seed):
    ops = _ops()
    from oracle import ldm_oracle as O
    r, g = _rng(400 + seed)
    off = r.choice([0, 0, 1, 2, 3])
    b, inner = r.randint(1, 7), r.choice([1, 3, 8, 250, 1024, r.randint(1, 9000)])
    sd = {"noise_scheduler.alpha_bar_t": O.noise_schedule(200)[2]}
    ab = sd["noise_scheduler.alpha_bar_t"].to(DEV)
    x0, eps = torch.randn((b, inner), generator=g), torch.randn((b, inner), generator=g)
    t = torch.tensor([r.randrange(200) for _ in range(b)])
    got = _guard(lambda: ops.q_sample(dev(x0, off), dev(eps, off), ab, t.to(DEV), lp=True), off, "q_sample")
    if got is not None:
        want = O.q_sample(sd, x0.view(b, inner, 1, 1), t, eps.view(b, inner, 1, 1)).view(b, inner)
        assert rel_err(got[0], want) <= 1e-6 and rel_err(got[1], want) <= 5e-3, f"q_sample b={b} inner={inner} off={off}"
        p = _guard(lambda: ops.predict_x0(got[0], dev(eps, off), ab, t.to(DEV)), off, "predict_x0")
        if p is not None:
            assert rel_err(p[0], O.predict_x0(sd, got[0].cpu().view(b, inner, 1, 1), t, eps.view(b, inner, 1, 1)).view(b, inner)) <= 1e-5
    n = r.choice([1, 5, 1003, 4096, r.randint(1, 50000)])
    p0, grads = torch.randn(n, generator=g), [torch.randn(n, generator=g) * (0.1 + i) for i in range(3)]
    refp = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([refp], lr=5e-4)
    p, m, v = p0.clone().to(DEV), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    for i, gr in enumerate(grads):
        refp.grad = gr.clone(); opt.step()
        ops.adam_step(p, gr.to(DEV), m, v, lr=5e-4, beta1=0.9, beta2=0.999, eps=1e-8, step=i + 1)
    assert rel_err(p, refp.detach()) <= 1e-6, f"adam n={n}"


@pytest.mark.parametrize("off", [1, 2, 3])
def test_vector_and_tma_entry_points_refuse_misaligned_tensors(off):
    """conv (tcgen05 / halo / edge / CUDA-core kernels), wgrad, split3, xattn, BatchNorm backward: a tensor that starts `off`
    elements into its allocation is refused with a message naming the alignment -- and the context is still healthy afterwards"""
    import math
    ops = _ops()
    from music_style_transfer_ldm_b200.ops import MstkError
    g = gen(900 + off)

    def refused(fn, what):
        with pytest.raises(MstkError, match="(?i)align"):
            fn()
        torch.cuda.synchronize()

    for dt, n, h, w, cin, cout in ((BF, 2, 16, 16, 64, 64), (F32, 2, 9, 7, 5, 6), (BF, 2, 16, 16, 1, 64), (F32, 1, 8, 8, 32, 16)):
        x = torch.randn((n, h, w, cin), generator=g).to(dt)
        wt = torch.randn((cout, cin, 3, 3), generator=g) / math.sqrt(9 * cin)
        wp = ops.pack_weight(wt.to(DEV), dt)
        refused(lambda: ops.conv(dev(x, off), wp, None, ksize=3), f"conv x {dt} cin={cin}")
        out = dev(torch.zeros((n, h, w, cout), dtype=dt), off)
        refused(lambda: ops.conv(dev(x, 0), wp, None, ksize=3, out=out), f"conv out {dt} cin={cin}")
        dy = torch.randn((n, h, w, cout), generator=g).to(dt)
        dw = torch.zeros(wt.shape, device=DEV)
        refused(lambda: ops.wgrad(dev(x, off), dev(dy, 0), dw, ksize=3), f"wgrad x {dt} cin={cin}")
        refused(lambda: ops.wgrad(dev(x, 0), dev(dy, off), dw, ksize=3), f"wgrad dy {dt} cin={cin}")
        y = ops.conv(dev(x, 0), wp, None, ksize=3, out_dtype=F32)           # and the aligned call still works
        ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.to(dt).float(), padding=1).permute(0, 2, 3, 1)
        assert rel_err(y, ref) <= 1e-4
    refused(lambda: ops.split3(dev(torch.randn((7, 64), generator=g), off)), "split3")
    v, dy = torch.randn((3, 4, 5, 16), generator=g), torch.randn((3, 4, 5, 16), generator=g)
    vec = lambda: torch.ones(16, device=DEV)
    refused(lambda: ops.bn_bwd(dev(dy, off), dev(v, 0), vec(), vec(), vec(), vec(), act=0, dgamma=vec(), dbeta=vec()), "bn_bwd")
