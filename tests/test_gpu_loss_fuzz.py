"""-m gpu: the public loss functions (music_style_transfer_ldm_b200/loss.py <- models/loss.py:10-101) on map sizes that are not
multiples of the VGGish stack's four 2x2 pools or of the kernels' tiles, and on odd batches, against the oracle.  The fused
kernels (first tap recomputed inside the statistics pass, taps fused with their pools) choose their tilings from the map size."""
import random

import pytest
import torch

from oracle import ldm_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _sizes(seed, count):
    rng = random.Random(seed)
    return [(rng.choice([1, 2, 3, 5]), rng.randint(16, 150), rng.randint(16, 200), rng.randrange(1 << 20)) for _ in range(count)]


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("bf16", 2e-2)])
@pytest.mark.parametrize("b,h,w,seed", _sizes(21, 8) + [(2, 16, 16, 1), (1, 17, 31, 2), (3, 128, 130, 3)], ids=lambda v: str(v))
def test_losses_at_any_size(sd, b, h, w, seed, precision, tol):
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200.loss import VGGishFeatureLoss
    g = torch.Generator().manual_seed(seed)
    x, y = torch.rand((b, 1, h, w), generator=g), torch.rand((b, 1, h, w), generator=g)
    z = torch.randn((b, 32, (h + 7) // 8, (w + 7) // 8), generator=g)
    z = z + 0.05 * torch.sign(z)
    fl = VGGishFeatureLoss()
    fl.load_state_dict({k[len("feature_loss_net."):]: v for k, v in sd.items() if k.startswith("feature_loss_net.")}, strict=True)
    fl.to(DEV)
    fl.precision = precision
    xd, yd, zd = x.to(DEV), y.to(DEV), z.to(DEV)
    assert float(fl(xd, yd)) == pytest.approx(float(O.vggish_feature_loss(sd, x, y)), rel=tol)
    assert float(M.style_loss(xd, yd, fl)) == pytest.approx(float(O.style_loss(sd, x, y)), rel=tol)
    assert float(M.compression_loss(xd, yd, zd, fl)) == pytest.approx(float(O.compression_loss(sd, x, y, z)), rel=tol)
    assert float(M.diffusion_loss(xd, yd)) == pytest.approx(float(O.diffusion_loss(x, y)), rel=1e-5)
    assert float(M.kl_regularization_loss(zd)) == pytest.approx(float(O.kl_regularization_loss(z)), rel=1e-5)
