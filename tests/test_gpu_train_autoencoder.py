"""-m gpu: the autoencoder stage (models/train.py:28-138) on libmstk kernels.
  * AutoencoderTrainer.train_step against tests/golden/ae_steps.npz (unmodified reference modules + AdamW): fp32 mode,
    losses of 3 consecutive steps within 1e-4 relative, updated parameters against the oracle trainer;
  * bf16 mode: 20-step loss curve against the fp32 oracle (first steps tight, 3 % on the chaotic tail -- see the test);
  * train_autoencoder(config, loaders): epoch loop, validation pass, scheduler, checkpoints.
"""
import numpy as np
import pytest
import torch

import cases
import make_golden_ae as GA

pytestmark = pytest.mark.gpu
DEV = "cuda"
# mathematically zero gradients: conv biases feeding a train-mode BatchNorm (see make_golden_ae.py)
NOISE = ("encoder.encoder.0.bias", "encoder.encoder.3.bias", "encoder.encoder.6.bias", "decoder.decoder.0.bias", "decoder.decoder.3.bias",
         "running_mean")


def _sub(sd, prefix):
    return {k[len(prefix):]: v.clone() for k, v in sd.items() if k.startswith(prefix)}


def _trainer(sd, precision, lr):
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200.loss import VGGishFeatureLoss
    from music_style_transfer_ldm_b200.train import AutoencoderTrainer
    enc, dec, fl = M.SpectrogramEncoder(32), M.SpectrogramDecoder(32), VGGishFeatureLoss()
    enc.load_state_dict(_sub(sd, "encoder."), strict=True)
    dec.load_state_dict(_sub(sd, "decoder."), strict=True)
    fl.load_state_dict(_sub(sd, "feature_loss_net."), strict=True)
    for m in (enc, dec, fl):
        m.to(DEV).set_precision(precision) if hasattr(m, "set_precision") else m.to(DEV)
    fl.precision = precision
    tr = AutoencoderTrainer(enc, dec, fl, torch.device(DEV), lr=lr)
    enc.train(); dec.train()
    return enc, dec, tr


@pytest.fixture
def cuda_core_fp32():
    """exact fp32 CUDA-core contractions (ops.FP32_TC = False) for the element-wise gradient comparison: the KL term's
    -1/z gradient is ill-conditioned (see above), so the ~1e-5 operand rounding of the default split-bf16 tensor-core
    route of fp32 mode is visible in it; the default route is held to the losses below"""
    from music_style_transfer_ldm_b200 import ops
    old, ops.FP32_TC = ops.FP32_TC, False
    yield
    ops.FP32_TC = old


def test_fp32_tensor_core_losses_match_reference_fixture(sd):
    from test_oracle_golden_train import AE_LOSS_RTOL
    g = cases.load("ae_steps")
    enc, dec, tr = _trainer(sd, "fp32", GA.LR)
    for i in range(GA.STEPS):
        got = tr.train_step(GA.step_input(i).to(DEV))
        np.testing.assert_allclose(got, g["losses"][i], rtol=max(AE_LOSS_RTOL[i], 2e-4), err_msg=f"step {i} loss vs reference modules + AdamW")


def test_fp32_steps_match_reference_fixture(sd, cuda_core_fp32):
    """losses of three steps and the step-0 gradients (read from the gradient arena) against the reference fixture;
    tolerances of the KL-conditioned encoder gradient as in tests/test_oracle_golden_train.py"""
    from test_oracle_golden_train import AE_LOSS_RTOL, ae_grad_tol
    g = cases.load("ae_steps")
    enc, dec, tr = _trainer(sd, "fp32", GA.LR)
    named = {**{"encoder." + k: p for k, p in enc.named_parameters()}, **{"decoder." + k: p for k, p in dec.named_parameters()}}
    for i in range(GA.STEPS):
        got = tr.train_step(GA.step_input(i).to(DEV))
        np.testing.assert_allclose(got, g["losses"][i], rtol=max(AE_LOSS_RTOL[i], 1e-4), err_msg=f"step {i} loss vs reference modules + AdamW")
        if i == 0:
            checked = 0
            for k in GA.SAVE_KEYS:
                if "g0:" + k in g:
                    want, have = g["g0:" + k], named[k].grad.detach().cpu().numpy()
                    assert np.abs(have - want).max() <= ae_grad_tol(k) * np.abs(want).max(), f"{k}: {np.abs(have - want).max()} of {np.abs(want).max()}"
                    checked += 1
            assert checked >= 8
    after = {k: p.detach().cpu() for k, p in named.items()}
    for k in GA.SAVE_KEYS:
        if k in after:                  # three Adam steps move a weight by at most ~3 lr either way (sign flips, see above)
            np.testing.assert_allclose(after[k].numpy(), g["p:" + k], rtol=1e-3, atol=7 * GA.LR, err_msg=k)
    np.testing.assert_allclose(enc.encoder[7].running_var.cpu().numpy(), g["p:encoder.encoder.7.running_var"], rtol=2e-3)
    assert int(enc.encoder[1].num_batches_tracked) == GA.STEPS and tr.step_count == GA.STEPS


def test_adamw_and_kl_grad_kernels():
    """mstk_adamw_step_dev against torch.optim.AdamW, mstk_kl_grad against autograd of loss.py:31-32"""
    from music_style_transfer_ldm_b200 import ops
    g = torch.Generator().manual_seed(2)
    p0, gr = torch.randn(1003, generator=g), torch.randn(1003, generator=g)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=1e-2, weight_decay=0.05)
    n = (1003 + 3) // 4 * 4
    p, gg, m, v = (torch.zeros(n, device=DEV) for _ in range(4))
    p[:1003] = p0.to(DEV)
    step = torch.zeros(1, dtype=torch.int64, device=DEV)
    for s in range(3):
        ref.grad = gr * (s + 1)
        opt.step()
        gg[:1003] = (gr * (s + 1)).to(DEV)
        step.add_(1)
        ops.adam_step(p, gg, m, v, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, step=step, weight_decay=0.05)
    np.testing.assert_allclose(p[:1003].cpu().numpy(), ref.detach().numpy(), rtol=2e-6, atol=1e-7)
    z = torch.randn((2, 16, 16, 32), generator=g)
    dz = torch.randn((2, 16, 16, 32), generator=g)
    zr = z.clone().requires_grad_(True)
    (0.01 * torch.mean(0.5 * (zr * zr - 1 - torch.log(zr * zr + 1e-8)))).backward()
    got = ops.kl_grad(z.to(DEV), dz.to(DEV), 0.01 / z.numel()).cpu()
    np.testing.assert_allclose(got.numpy(), (zr.grad + dz).numpy(), rtol=1e-5, atol=1e-7)
    got = ops.kl_grad(z.to(DEV), None, 0.01 / z.numel()).cpu()
    np.testing.assert_allclose(got.numpy(), zr.grad.numpy(), rtol=1e-5, atol=1e-9)


def test_bf16_loss_curve_within_one_percent(sd):
    from oracle import train_oracle as T
    enc, dec, tr = _trainer(sd, "bf16", GA.LR)
    ot = T.OracleAutoencoderTrainer(sd, lr=GA.LR)
    for i in range(20):
        x = torch.rand((4, 1, 128, 128), generator=torch.Generator().manual_seed(9100 + i))
        got, want = tr.train_step(x.to(DEV)), ot.train_step(x)
        # The KL term's -1/z gradient (make_golden_ae.py) makes the TRAJECTORY chaotic in the reference itself: fp32 runs that
        # differ only in summation order are 0.6 % apart by step 11 (tools/ae_variability.py: max over 6 runs x 20 steps =
        # 0.62 % fp32, 0.94 % bf16, not growing).  Tight on the first steps, 3 % on the rest.
        tol = (2e-3, 5e-3)[i] if i < 2 else 3e-2
        assert abs(got - want) <= tol * abs(want), f"step {i}: {got} vs oracle {want}"


def test_graph_replayed_step_matches_eager(sd):
    enc1, dec1, tr1 = _trainer(sd, "fp32", GA.LR)
    enc2, dec2, tr2 = _trainer(sd, "fp32", GA.LR)
    tr2.use_cuda_graph = False
    for i in range(5):
        x = torch.rand((4, 1, 128, 128), generator=torch.Generator().manual_seed(9300 + i)).to(DEV)
        a, b = tr1.train_step(x), tr2.train_step(x)
        # two runs of the SAME code differ through atomic summation order amplified by the KL gradient (see above)
        assert abs(a - b) <= (1e-4 if i < 2 else 2e-2) * abs(b), f"step {i}: graph {a} vs eager {b}"
        if i == 3:
            tr1.lr = tr2.lr = GA.LR / 2                                                # re-capture on a learning-rate change
    assert tr1._graph is not None and tr2._graph is None and tr1.step_count == tr2.step_count == 5
    assert int(enc1.encoder[1].num_batches_tracked) == 5


def test_errors(sd):
    from music_style_transfer_ldm_b200.ops import MstkError
    enc, dec, tr = _trainer(sd, "fp32", 1e-3)
    with pytest.raises(MstkError):
        tr.train_step(torch.rand(2, 1, 128, 128))                   # CPU tensor: no fallback
    with pytest.raises(MstkError):
        tr.train_step(torch.rand(2, 1, 100, 128, device=DEV))
    enc.eval()
    with pytest.raises(MstkError):
        tr.train_step(torch.rand(2, 1, 128, 128, device=DEV))


def test_train_autoencoder_loop(tmp_path, capsys):
    from music_style_transfer_ldm_b200.config import config
    from music_style_transfer_ldm_b200.train import train_autoencoder
    g = torch.Generator().manual_seed(1)
    base = torch.rand((8, 1, 128, 128), generator=g)
    train_loader = [(base[:4],), (base[4:],)] * 3
    test_loader = [(base[:4],)]
    cfg = dict(config, num_epochs=3, learning_rate=2e-3)
    torch.manual_seed(0)
    enc, dec, tl, vl = train_autoencoder(cfg, train_loader, test_loader, checkpoint_dir=str(tmp_path))
    assert len(tl) == 3 and len(vl) == 3 and np.isfinite(tl + vl).all()
    assert tl[-1] < tl[0]                                           # it learns the 8 fixed images
    assert (tmp_path / "encoder.pth").is_file() and (tmp_path / "decoder.pth").is_file()
    assert set(torch.load(tmp_path / "encoder.pth", map_location="cpu")) == set(enc.state_dict())
    assert "Average Val Loss" in capsys.readouterr().out
    with pytest.raises(Exception):
        train_autoencoder(cfg)
