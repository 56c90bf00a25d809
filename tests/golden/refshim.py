"""Import shim for the UNMODIFIED reference modules (build container only).

/root/reference does not exist on the GPU box, so nothing in the ``-m gpu``
tests, smoke() or bench.py may import this file.  It is used by
``make_golden.py`` (fixture generation) and by ``tests/test_reference_live.py``
(skipped automatically when the reference tree is absent).

What is shimmed, and why:
  * pytorch_lightning, resampy, soundfile, matplotlib: imported but unused on
    the hot path and not installed here -> empty stub modules.
  * lpips: not installed, weights need network -> stub; the config switch
    ``compression_feature_extractor`` is set to the reference's other supported
    branch, 'vggish' (loss.py:25-28).
  * torch.hub.load('harritaylor/torchvggish', 'vggish') (loss.py:56) needs the
    network -> returns the vendored architecture (thirdparty/torchvggish/
    vggish.py:108-118) with random weights.
  * train.py:22-26 raises without CUDA -> torch.cuda.is_available patched for the
    duration of ``import train`` only.
"""
from __future__ import annotations

import os
import sys
import types

import torch

REFERENCE_ROOT = os.environ.get("MSTK_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "models", "model.py"))


_cache = {}


def load(with_train: bool = False):
    """Returns a namespace with .model, .loss, .config (and .train) reference modules."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    if "ns" in _cache and (not with_train or hasattr(_cache["ns"], "train")):
        return _cache["ns"]
    for name in ("pytorch_lightning", "resampy", "soundfile", "tqdm_stub"):
        sys.modules.setdefault(name, types.ModuleType(name))
    lp = types.ModuleType("lpips")
    lp.LPIPS = None
    sys.modules["lpips"] = lp
    if "matplotlib" not in sys.modules:
        mpl, plt = types.ModuleType("matplotlib"), types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, plt
    for p in (os.path.join(REFERENCE_ROOT, "thirdparty"), os.path.join(REFERENCE_ROOT, "models")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from torchvggish.vggish import make_layers  # vendored architecture only

    class _VGGishFeaturesOnly(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.features = make_layers()

    torch.hub.load = lambda *a, **k: _VGGishFeaturesOnly()
    import config as ref_config
    ref_config.config["compression_feature_extractor"] = "vggish"
    import loss as ref_loss
    import model as ref_model
    ns = types.SimpleNamespace(model=ref_model, loss=ref_loss, config=ref_config)
    if with_train:
        real = torch.cuda.is_available
        torch.cuda.is_available = lambda: True
        try:
            import train as ref_train
        finally:
            torch.cuda.is_available = real
        ns.train = ref_train
    _cache["ns"] = ns
    return ns


def build_ldm(sd, latent_dim: int = 32):
    """Reference LDM (pretrained_path=None skips .pth loading, model.py:262) carrying ``sd``."""
    ns = load()
    ldm = ns.model.LDM(latent_dim=latent_dim, pretrained_path=None)
    missing, unexpected = ldm.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    return ldm


class BatchedSamplingLDM:
    """Context manager that neutralises the per-step ``t.item()`` logging line
    (model.py:461, 555), which raises for batch > 1.  The arithmetic is untouched:
    inside the context Tensor.item() on a multi-element tensor returns element 0."""

    def __enter__(self):
        self._real = torch.Tensor.item

        def item(t):
            return self._real(t.reshape(-1)[0]) if t.numel() > 1 else self._real(t)

        torch.Tensor.item = item
        return self

    def __exit__(self, *exc):
        torch.Tensor.item = self._real
        return False
