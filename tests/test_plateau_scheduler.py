"""CPU: PlateauScheduler (host logic of LDMTrainer.train / train_autoencoder) against torch's ReduceLROnPlateau,
the scheduler the reference constructs at train.py:48 and train.py:159-161."""
import pytest
import torch

from music_style_transfer_ldm_b200.train import PlateauScheduler


class _Owner:
    def __init__(self, lr):
        self.lr = lr


@pytest.mark.parametrize("factor,patience,min_lr", [(0.5, 10, 0.0), (0.5, 5, 1e-6), (0.1, 0, 1e-5)])
def test_matches_torch_reduce_lr_on_plateau(factor, patience, min_lr):
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([p], lr=5e-4)
    ref = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=factor, patience=patience, min_lr=min_lr)
    own = _Owner(5e-4)
    mine = PlateauScheduler(own, mode_factor=factor, patience=patience, min_lr=min_lr)
    g = torch.Generator().manual_seed(3)
    level = 1.0
    for i in range(300):
        if i % 37 == 0:
            level *= 0.8                                           # occasional real improvements between plateaus
        metric = level + 0.01 * float(torch.rand((), generator=g))
        ref.step(metric)
        mine.step(metric)
        assert own.lr == pytest.approx(opt.param_groups[0]["lr"], rel=1e-12), f"epoch {i}"
    assert own.lr < 5e-4


def test_rejects_factor_of_one():
    with pytest.raises(ValueError):
        PlateauScheduler(_Owner(1e-3), mode_factor=1.0)
