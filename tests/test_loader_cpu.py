"""CPU: LDM.__init__'s weight-loading branches (reference models/model.py:262-353) -- M13 of SURVEY 8(a).

  * ``encoder.pth`` / ``decoder.pth`` branch: VAE loaded, encoder frozen + eval, decoder trainable + train, and the
    reference's quirk of RE-CREATING unet / noise scheduler / style encoder afterwards (model.py:350-353);
  * ``load_full_model=True``: prefix-split ``ldm.pth`` into the five components and return early (model.py:262-315);
  * the fallback from a missing ``ldm.pth`` to the VAE-only branch (model.py:316-318);
  * LIVE against the unmodified reference (only where /root/reference exists, i.e. the build container): same seed ->
    identical state_dict, requires_grad flags and module modes for every branch.
"""
import os
import sys

import pytest
import torch

import cases

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import refshim  # noqa: E402


@pytest.fixture()
def ckpt_dir(tmp_path, sd):
    d = str(tmp_path) + os.sep
    torch.save({k[len("encoder."):]: v for k, v in sd.items() if k.startswith("encoder.")}, d + "encoder.pth")
    torch.save({k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")}, d + "decoder.pth")
    return d


def _ldm(*a, **kw):
    import music_style_transfer_ldm_b200 as M
    return M.LDM(*a, **kw)


def _same(a, b):
    return a.shape == b.shape and torch.equal(a.detach().cpu(), b.detach().cpu())


def test_vae_only_branch(ckpt_dir, sd):
    m = _ldm(latent_dim=32, pretrained_path=ckpt_dir)
    cur = m.state_dict()
    assert all(_same(cur[k], v) for k, v in sd.items() if k.startswith(("encoder.", "decoder.")))
    assert not any(p.requires_grad for p in m.encoder.parameters()) and all(p.requires_grad for p in m.decoder.parameters())
    assert not m.encoder.training and m.decoder.training and m.style_encoder.training
    # unet / style encoder are re-created after loading: fresh initialisation, not the checkpoint's values
    assert not _same(cur["unet.enc1.weight"], sd["unet.enc1.weight"])
    assert all(p.requires_grad for p in m.unet.parameters())


def test_full_model_branch_and_fallback(ckpt_dir, sd):
    torch.save(dict(sd), ckpt_dir + "ldm.pth")
    m = _ldm(latent_dim=32, pretrained_path=ckpt_dir, load_full_model=True)
    cur = m.state_dict()
    for k, v in sd.items():
        if k.startswith(("encoder.", "decoder.", "unet.", "style_encoder.", "noise_scheduler.")):
            assert _same(cur[k], v), k
    assert not any(p.requires_grad for p in m.encoder.parameters())
    assert not m.encoder.training and m.decoder.training and m.unet.training and m.style_encoder.training
    os.remove(ckpt_dir + "ldm.pth")
    m2 = _ldm(latent_dim=32, pretrained_path=ckpt_dir, load_full_model=True)          # model.py:316-318
    assert _same(m2.state_dict()["decoder.decoder.0.weight"], sd["decoder.decoder.0.weight"])
    assert not _same(m2.state_dict()["unet.enc1.weight"], sd["unet.enc1.weight"])
    with pytest.raises(FileNotFoundError):
        _ldm(latent_dim=32, pretrained_path=ckpt_dir + "missing" + os.sep)


@pytest.mark.skipif(not refshim.available(), reason="the reference tree exists only in the build container")
@pytest.mark.parametrize("branch", ["none", "vae", "full"])
def test_constructor_matches_live_reference(ckpt_dir, sd, branch):
    ref = refshim.load().model
    if branch == "full":
        torch.save(dict(sd), ckpt_dir + "ldm.pth")
    kw = {"none": dict(pretrained_path=None), "vae": dict(pretrained_path=ckpt_dir),
          "full": dict(pretrained_path=ckpt_dir, load_full_model=True)}[branch]
    torch.manual_seed(123)
    r = ref.LDM(latent_dim=32, **kw)
    torch.manual_seed(123)
    m = _ldm(latent_dim=32, **kw)
    rs, ms = r.state_dict(), m.state_dict()
    keys = [k for k in rs if not k.startswith("feature_loss_net.")]
    assert set(keys) == {k for k in ms if not k.startswith("feature_loss_net.")}
    for k in keys:
        assert _same(rs[k], ms[k]), f"{branch}: {k} differs from the reference constructor under the same seed"
    rp, mp = dict(r.named_parameters()), dict(m.named_parameters())
    for k in rp:
        if not k.startswith("feature_loss_net."):
            assert rp[k].requires_grad == mp[k].requires_grad, k
    for name in ("encoder", "decoder", "unet", "style_encoder"):
        assert getattr(r, name).training == getattr(m, name).training, name
