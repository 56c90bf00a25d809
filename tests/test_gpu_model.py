"""-m gpu: the drop-in modules (CUDA path through the C ABI) against the golden fixtures that
tests/golden/make_golden.py produced from the UNMODIFIED reference, and against the oracle.
fp32 mode: 1e-4 relative (max norm); bf16 mode: 2e-2 relative."""
import numpy as np
import pytest
import torch

import cases
from gpu_util import TOL, assert_close, assert_close_mode, rel_err, rel_l2

pytestmark = pytest.mark.gpu
DEV = "cuda"
PRECISIONS = ["fp32", "bf16"]


@pytest.fixture(scope="module")
def ldm_by_precision(sd):
    import music_style_transfer_ldm_b200 as M
    out = {}
    for p in PRECISIONS:
        m = M.LDM(latent_dim=32, pretrained_path=None)
        m.load_state_dict(sd, strict=True)
        m.to(DEV).train()          # the reference's scripts leave a fresh LDM in train mode
        m.set_precision(p)
        out[p] = m
    return out


def _reset(m, sd):
    m.load_state_dict(sd, strict=True)   # BatchNorm running buffers are updated in train mode


class fixed_randn_like:
    def __init__(self, eps):
        self.eps = eps

    def __enter__(self):
        self._real = torch.randn_like
        torch.randn_like = lambda x, **kw: self.eps.to(device=x.device, dtype=x.dtype)
        return self

    def __exit__(self, *exc):
        torch.randn_like = self._real
        return False


class fixed_randn:
    def __init__(self, z):
        self.z = z

    def __enter__(self):
        self._real = torch.randn
        torch.randn = lambda *a, **kw: self.z.clone()
        return self

    def __exit__(self, *exc):
        torch.randn = self._real
        return False


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name", ["unet_c1", "unet_b3", "unet_wide"])
def test_unet_golden(ldm_by_precision, precision, name):
    m = ldm_by_precision[precision]
    c, g = cases.CASES[name], cases.load(name)
    h, w = c.get("h", 16), c.get("w", 16)
    z, st = cases.latent(11, c["b"], h, w), cases.style_dict(12, c["b"], h, w)
    out = m.unet(z.to(DEV), torch.tensor(c["t"]).to(DEV), {k: v.to(DEV) for k, v in st.items()})
    assert out.shape == z.shape and out.dtype == torch.float32
    assert_close_mode(out, g["eps_hat"], precision, name)


@pytest.mark.parametrize("precision", PRECISIONS)
def test_unet_float_timestep_and_layers(ldm_by_precision, precision, sd):
    m = ldm_by_precision[precision]
    g = cases.load("unet_float_t")
    z, st = cases.latent(11, 1), cases.style_dict(12, 1)
    out = m.unet(z.to(DEV), torch.tensor([37.0]).to(DEV), {k: v.to(DEV) for k, v in st.items()})
    assert_close_mode(out, g["eps_hat"], precision, "float t")
    # layer-by-layer against the oracle's taps (localises a failing kernel)
    from oracle import ldm_oracle as O
    from music_style_transfer_ldm_b200 import ops
    from music_style_transfer_ldm_b200.model import _act_dtype
    adt = _act_dtype(precision)
    z, st = cases.latent(11, 3), cases.style_dict(12, 3)
    t = torch.tensor([0, 57, 199])
    ref_taps = {}
    O.unet_fwd(sd, z, t, st, taps=ref_taps)
    taps = {}
    u = m.unet
    kv2, kv1 = u._project_style(ops.nchw_to_nhwc(st["s5"].to(DEV), adt), ops.nchw_to_nhwc(st["s6"].to(DEV), adt))
    u._forward_nhwc(ops.nchw_to_nhwc(z.to(DEV), adt), u._time_embedding(t.to(DEV)), kv2, kv1, taps=taps)
    for k in ("z1", "z2", "z3", "a3", "z4", "a4", "bt", "d4", "d3", "d2"):
        assert_close_mode(ops.nhwc_to_nchw(taps[k]), ref_taps[k], precision, f"tap {k}")


@pytest.mark.parametrize("precision", PRECISIONS)
def test_style_encoder_golden(ldm_by_precision, precision):
    m = ldm_by_precision[precision]
    g = cases.load("style_enc")
    emb = m.style_encoder(cases.spec(21, 2).to(DEV))
    assert list(emb.keys()) == ["s1", "s2", "s3", "s4", "s5", "s6"]
    assert emb["s1"].shape == (2, 64, 64, 64) and emb["s6"].shape == (2, 512, 2, 2)
    assert_close_mode(emb["s5"], g["s5"], precision, "s5")
    assert_close_mode(emb["s6"], g["s6"], precision, "s6")
    for i in range(6):
        got = emb[f"s{i + 1}"].double().sum().item()
        assert got == pytest.approx(g["tap_sums"][i], rel=10 * TOL[precision])


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name", ["vae_train", "vae_eval", "vae_wide_eval"])
def test_vae_golden(ldm_by_precision, precision, name, sd):
    m = ldm_by_precision[precision]
    c, g = cases.CASES[name], cases.load(name)
    x = cases.spec(31, c["b"], c.get("h", 128), c.get("w", 128)).to(DEV)
    m.encoder.train(c["training"]); m.decoder.train(c["training"])
    try:
        z0 = m.encoder(x)
        y = m.decoder(z0)
        assert z0.shape == g["z0"].shape and y.shape == g["decoded"].shape
        assert_close_mode(z0, g["z0"], precision, "z0")
        assert_close_mode(y, g["decoded"], precision, "decoded")
        assert y.abs().max().item() <= 1.0
        if c["training"]:
            # running buffers moved exactly as nn.BatchNorm2d moves them
            from oracle import ldm_oracle as O
            upd = {}
            O.decoder_fwd(sd, O.encoder_fwd(sd, x.cpu(), True, update=upd), True, update=upd)
            cur = m.state_dict()
            for k, v in upd.items():
                if k.endswith("num_batches_tracked"):
                    assert int(cur[k]) == int(v)
                else:
                    assert_close(cur[k], v, 10 * TOL[precision], k)
    finally:
        _reset(m, sd)
        m.train()


@pytest.mark.parametrize("precision", PRECISIONS)
def test_ldm_forward_and_losses_golden(ldm_by_precision, precision, sd):
    import music_style_transfer_ldm_b200 as M
    m = ldm_by_precision[precision]
    c, g, gl = cases.CASES["ldm_forward"], cases.load("ldm_forward"), cases.load("losses")
    x, s, eps = cases.spec(41, c["b"]), cases.spec(42, c["b"]), cases.latent(43, c["b"])
    try:
        with fixed_randn_like(eps):
            out = m(x.to(DEV), s.to(DEV), torch.tensor(c["t"]).to(DEV))
        assert set(out) == {"z_t", "noise", "noise_pred", "z_0", "reconstructed"}
        assert torch.equal(out["noise"].cpu(), eps)
        for k in ("z_t", "noise_pred", "z_0", "reconstructed"):
            assert_close_mode(out[k], g[k], precision, k)
        tol = 10 * TOL[precision]
        den = M.diffusion_loss(out["noise_pred"], out["noise"])
        kl = M.kl_regularization_loss(out["z_0"])
        comp = M.compression_loss(x.to(DEV), out["reconstructed"], out["z_0"], m.feature_loss_net)
        sty = M.style_loss(out["reconstructed"], s.to(DEV), m.feature_loss_net)
        assert den.dim() == 0
        assert den.item() == pytest.approx(float(gl["diffusion"]), rel=tol)
        assert kl.item() == pytest.approx(float(gl["kl"]), rel=tol)
        assert comp.item() == pytest.approx(float(gl["compression"]), rel=tol)
        assert sty.item() == pytest.approx(float(gl["style"]), rel=tol)
        assert m.feature_loss_net(x.to(DEV), s.to(DEV)).item() == pytest.approx(float(gl["vggish_x_s"]), rel=tol)
        assert (comp + den + 3.0 * sty).item() == pytest.approx(float(gl["total"]), rel=tol)
    finally:
        _reset(m, sd)


@pytest.mark.parametrize("graphs", [False, True])
@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name", ["transfer_b1", "transfer_b2"])
def test_transfer_golden(ldm_by_precision, precision, name, graphs, sd):
    m = ldm_by_precision[precision]
    c, g = cases.CASES[name], cases.load(name)
    x, s, eps = cases.spec(51, c["b"]), cases.spec(52, c["b"]), cases.latent(53, c["b"])
    m.use_cuda_graphs = graphs
    try:
        for _ in range(2 if graphs else 1):      # second call replays the cached graph
            _reset(m, sd)
            with fixed_randn_like(eps):
                dec, ztd = m.content_style_transfer_wrapper(x.to(DEV), s.to(DEV), num_timesteps=c["n"], eta=c["eta"])
            assert dec.shape == (c["b"], 1, 128, 128) and ztd.shape == (c["b"], 1, 128, 128)
            assert_close_mode(dec, g["decoded"], precision, "decoded")
            assert_close_mode(ztd, g["z_t_decoded"], precision, "z_t_decoded")
    finally:
        m.use_cuda_graphs = True
        _reset(m, sd)


@pytest.mark.parametrize("precision", PRECISIONS)
def test_ddim_inner_logs_golden(ldm_by_precision, precision):
    m = ldm_by_precision[precision]
    g = cases.load("ddim_inner")
    z_t, st = cases.latent(54, 1), cases.style_dict(55, 1)
    m.collect_sampling_logs = True
    try:
        x, logs = m.content_style_ddim_sample(z_t.to(DEV), {k: v.to(DEV) for k, v in st.items()}, timesteps=5, eta=0.25)
    finally:
        m.collect_sampling_logs = False
    assert logs["timesteps"] == g["timesteps"].tolist() == [4, 3, 2, 1]      # N-1 UNet calls
    assert_close_mode(x, g["x"], precision, "x")
    assert_close_mode(torch.stack(logs["pred_x0"]), g["pred_x0"], precision, "pred_x0")
    assert_close_mode(torch.stack(logs["noise_pred"]), g["noise_pred"], precision, "noise_pred")
    x2, logs2 = m.content_style_ddim_sample(z_t.to(DEV), {k: v.to(DEV) for k, v in st.items()}, timesteps=5, eta=0.25)
    assert logs2["timesteps"] == [4, 3, 2, 1] and logs2["pred_x0"] == []
    assert_close_mode(x2, g["x"], precision, "x (no logs)")


@pytest.mark.parametrize("precision", PRECISIONS)
def test_style_only_golden(ldm_by_precision, precision, sd):
    m = ldm_by_precision[precision]
    c, g = cases.CASES["style_only_b1"], cases.load("style_only_b1")
    s, zT = cases.spec(61, c["b"]), cases.latent(62, c["b"])
    try:
        with fixed_randn(zT):
            dec = m.style_ddim_sample_wrapper((c["b"], 32, 16, 16), s.to(DEV), timesteps=c["n"], eta=c["eta"])
        assert_close_mode(dec, g["decoded"], precision, "decoded")
    finally:
        _reset(m, sd)


def test_full_size_properties(ldm_by_precision, sd):
    """BASELINE config-3 size (batch 64, 50 steps): size-independent properties instead of an
    oracle run -- batch rows are independent in eval mode, so sample i of the big batch must
    equal the same sample run alone; and graph replay is bit-identical to eager."""
    m = ldm_by_precision["bf16"]
    m.eval()
    try:
        x, s, eps = cases.spec(71, 64), cases.spec(72, 64), cases.latent(73, 64)
        with fixed_randn_like(eps):
            dec, ztd = m.content_style_transfer_wrapper(x.to(DEV), s.to(DEV), num_timesteps=50, eta=0.0)
        assert torch.isfinite(dec).all() and dec.min() >= 0 and dec.max() <= 1
        with fixed_randn_like(eps[5:7]):
            dec2, _ = m.content_style_transfer_wrapper(x[5:7].to(DEV), s[5:7].to(DEV), num_timesteps=50, eta=0.0)
        assert rel_l2(dec[5:7], dec2) < 2e-2
        m.use_cuda_graphs = False
        with fixed_randn_like(eps):
            dec3, _ = m.content_style_transfer_wrapper(x.to(DEV), s.to(DEV), num_timesteps=50, eta=0.0)
        assert rel_l2(dec3, dec) < 2e-2
    finally:
        m.use_cuda_graphs = True
        m.train()
        _reset(m, sd)


def test_transfer_batches_equals_the_wrapper_call_by_call(ldm_by_precision, sd):
    """LDM.transfer_batches (uploads / downloads pipelined on copy streams around the whole-call graph): every yielded pair equals
    the wrapper called on the same batch with the same generator state, in order, as pinned CPU tensors (fp32 mode, 1e-5:
    split-K partial sums are reduced in arrival order, so two runs of the SAME call already differ in the last bits)"""
    m = ldm_by_precision["fp32"]
    m.eval()
    try:
        g = torch.Generator().manual_seed(81)
        batches = [(torch.rand((3, 1, 128, 128), generator=g).pin_memory(), torch.rand((3, 1, 128, 128), generator=g)) for _ in range(4)]
        torch.manual_seed(5); torch.cuda.manual_seed(5)
        want = [tuple(t.cpu() for t in m.content_style_transfer_wrapper(c.to(DEV), s.to(DEV), num_timesteps=6, eta=0.0)) for c, s in batches]
        torch.manual_seed(5); torch.cuda.manual_seed(5)
        got = list(m.transfer_batches(iter(batches), num_timesteps=6, eta=0.0))
        assert len(got) == 4
        for (d, z), (dw, zw) in zip(got, want):
            assert not d.is_cuda and d.is_pinned() and d.shape == (3, 1, 128, 128)
            assert rel_err(d, dw) <= 1e-5 and rel_err(z, zw) <= 1e-5, (rel_err(d, dw), rel_err(z, zw))
        assert list(m.transfer_batches([], num_timesteps=6)) == []
        # a ring of pre-allocated pinned pairs, results one call late: the consumer copies before asking for the next
        torch.manual_seed(5); torch.cuda.manual_seed(5)
        ringed = [(d.clone(), z.clone()) for d, z in m.transfer_batches(iter(batches), num_timesteps=6, eta=0.0, ring=3, depth=1)]
        assert len(ringed) == 4 and all(rel_err(d, dw) <= 1e-5 and rel_err(z, zw) <= 1e-5 for (d, z), (dw, zw) in zip(ringed, want))
        from music_style_transfer_ldm_b200.ops import MstkError
        with pytest.raises(MstkError, match="ring"):
            list(m.transfer_batches(iter(batches), num_timesteps=6, ring=3, depth=2))
    finally:
        m.train()
        _reset(m, sd)
