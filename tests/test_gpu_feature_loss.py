"""-m gpu: the vectorised VGGish feature-loss kernels (csrc/feature_loss.cu) against the reference formula
(models/loss.py:94-99: x / (std over (C,H,W), unbiased, + 1e-8), MSE) computed with torch on the CPU."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import assert_close, gen, nchw, nhwc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _term(f, h):
    fn = f / (f.std(dim=[1, 2, 3], keepdim=True) + 1e-8)
    hn = h / (h.std(dim=[1, 2, 3], keepdim=True) + 1e-8)
    return ((fn - hn) ** 2).mean().reshape(1)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(3, 64, 16, 16), (5, 40, 9, 8), (2, 512, 8, 8)])
def test_feature_mse3(shape, dt):
    from music_style_transfer_ldm_b200 import ops
    g = gen(sum(shape))
    b = shape[0]
    maps = [(torch.relu(torch.randn(shape, generator=g)) * s).to(dt).float() for s in (1.0, 3.0, 0.25)]
    ref_a, ref_c = _term(maps[0], maps[1]), _term(maps[1], maps[2])
    f_all = torch.cat([nhwc(m, dt) for m in maps], dim=0).to(DEV)
    for _ in range(2):      # twice: statistics and scratch must come back zeroed
        out_a, out_c = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
        ops.feature_mse3(f_all, b, 0.5, out_a, out_c)
        ops.feature_mse3(f_all, b, 0.5, out_a, out_c)
        assert_close(out_a, ref_a, 2e-5, "feature_mse3 pair (0,1)")
        assert_close(out_c, ref_c, 2e-5, "feature_mse3 pair (1,2)")
    # the two-map entry point goes through the same kernels
    out = torch.zeros(1, device=DEV)
    ops.feature_std_mse(f_all[:b], f_all[b:2 * b], 1.0, out)
    assert_close(out, ref_a, 2e-5, "feature_std_mse (vector path)")


def test_feature_std_mse_scalar_fallback():
    from music_style_transfer_ldm_b200 import ops
    g = gen(3)
    f, h = torch.relu(torch.randn((3, 5, 3, 7), generator=g)), torch.relu(torch.randn((3, 5, 3, 7), generator=g)) * 2    # inner = 105: odd
    out = torch.zeros(1, device=DEV)
    ops.feature_std_mse(nhwc(f).to(DEV), nhwc(h).to(DEV), 1.0, out)
    assert_close(out, _term(f, h), 1e-5, "feature_std_mse (scalar path)")


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(3, 64, 16, 32), (2, 24, 8, 12), (2, 20, 6, 10)])
def test_maxpool_vec(shape, dt):
    from music_style_transfer_ldm_b200 import ops
    x = torch.randn(shape, generator=gen(shape[1])).to(dt).float()
    got = nchw(ops.maxpool2x2(nhwc(x, dt).to(DEV))).cpu()
    assert torch.equal(got, F.max_pool2d(x, 2, 2))


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape,pool", [((3, 64, 16, 16), True), ((2, 128, 8, 12), False), ((4, 8, 6, 10), True)])
def test_feature_fused3(shape, pool, dt):
    """one-pass tap (sums + cross terms + max pool) vs the two-pass reference formula"""
    from music_style_transfer_ldm_b200 import ops
    g = gen(sum(shape) + 1)
    b = shape[0]
    maps = [(torch.relu(torch.randn(shape, generator=g)) * s).to(dt).float() for s in (1.0, 3.0, 0.25)]
    # make the first pair strongly correlated: the algebraic form must survive the cancellation
    maps[1] = (maps[0] * 2.0 + 0.05 * torch.relu(torch.randn(shape, generator=g))).to(dt).float()
    ref_a, ref_c = _term(maps[0], maps[1]), _term(maps[1], maps[2])
    f_all = torch.cat([nhwc(m, dt) for m in maps], dim=0).to(DEV)
    for _ in range(2):      # accumulators and counter must come back zeroed
        out_a, out_c = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
        pooled = ops.feature_fused3(f_all, b, 1.0, out_a, out_c, pool)
        assert_close(out_a, ref_a, 1e-3, "feature_fused3 pair (0,1)")
        assert_close(out_c, ref_c, 1e-4, "feature_fused3 pair (1,2)")
        if pool:
            want = F.max_pool2d(torch.cat(maps, dim=0), 2, 2)
            assert torch.equal(nchw(pooled).cpu(), want)
        else:
            assert pooled is None


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("b,h,w", [(2, 128, 128), (3, 16, 24)])
def test_feature_first3_never_materialises_the_first_tap(b, h, w, dt):
    """conv(1 -> 64) + ReLU recomputed inside the statistics pass: the two loss terms of the tap against the reference
    formula on the fp32 tap, the pooled map against max_pool2d of the tap"""
    from music_style_transfer_ldm_b200 import ops
    g = gen(b + h)
    imgs = [torch.rand((b, 1, h, w), generator=g) for _ in range(3)]
    imgs[1] = (imgs[0] * 0.9 + 0.1 * imgs[1]).contiguous()
    wt = torch.randn((64, 1, 3, 3), generator=g) * 0.3
    bias = torch.randn((64,), generator=g) * 0.1
    taps = [F.relu(F.conv2d(im, wt, bias, padding=1)) for im in imgs]
    ref_a, ref_c = _term(taps[0], taps[1]), _term(taps[1], taps[2])
    three = torch.cat(imgs, dim=0).view(3 * b, h, w, 1).to(DEV)
    wp = ops.pack_weight(wt.to(DEV), torch.float32)
    for _ in range(2):
        out_a, out_c = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
        pooled = ops.feature_first3(three, wp, bias.to(DEV), b, 1.0, out_a, out_c, dt)
        assert_close(out_a, ref_a, 1e-4, "first tap, pair (x, recon)")
        assert_close(out_c, ref_c, 1e-4, "first tap, pair (recon, style)")
        want = F.max_pool2d(torch.cat(taps, dim=0), 2, 2)
        assert pooled.dtype == dt
        # bf16 map: the tensor-core kernel carries pixels and weights as bf16 hi + lo (~1e-5 relative before rounding), so a value
        # on a bf16 rounding boundary may land one ulp (2^-7 of the value) away from the rounded fp32 tap -- on very few elements
        assert_close(nchw(pooled), want.to(dt).float(), 1e-6 if dt == torch.float32 else 8e-3, "pooled first tap")
        if dt == torch.bfloat16:
            frac = (nchw(pooled).cpu() != want.to(dt).float()).float().mean().item()
            assert frac < 5e-3, f"{frac:.3%} of the pooled bf16 values differ from the rounded fp32 tap"
