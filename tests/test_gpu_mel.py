"""Mel front end, GPU side: csrc/mel.cu through the C ABI vs the numpy oracle (librosa restatement)."""
import numpy as np
import pytest
import torch

from oracle import mel_oracle as MO

pytestmark = pytest.mark.gpu


def _waves(b, n, seed=0):
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(n) / 22050.0
    out = []
    for i in range(b):
        f0 = 110.0 * (i + 1)
        y = 0.3 * torch.sin(2 * np.pi * f0 * t) + 0.1 * torch.sin(2 * np.pi * 7 * f0 * t) * torch.exp(-2 * t)
        out.append(y + (0.003 * (i + 1)) * torch.randn(n, generator=g))
    return torch.stack(out).float()


@pytest.mark.parametrize("b,n,crop", [(3, 66150, 128), (2, 220500, 384), (1, 5000, 9)])
def test_mel_frontend_vs_oracle(b, n, crop):
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200.audio_processor import AudioPreprocessor
    ap = AudioPreprocessor()
    w = _waves(b, n)
    db = ap.get_mel_spectogram(w.cuda(), 22050, n_mels=128).cpu().numpy()
    img = ap.spectrogram_batch(w.cuda(), n_mels=128, crop_frames=crop).cpu().numpy()
    frames = 1 + n // 512
    assert db.shape == (b, 128, frames) and img.shape == (b, 1, 128, crop)
    for i in range(b):
        ref_db = MO.get_mel_spectogram(w[i].numpy(), 22050, 128)
        # fp32 FFT on both sides; bins near the -80 dB floor are 8 orders of magnitude below the peak,
        # so compare in dB with an absolute tolerance (0.02 dB = 0.06 grey levels)
        np.testing.assert_allclose(db[i], ref_db, atol=2e-2)
        ref_img = MO.to_unit_tensor(MO.to_grayscale_u8(ref_db), crop)
        lv = np.abs(np.rint(img[i, 0] * 255) - np.rint(ref_img * 255))
        assert lv.max() <= 1, "uint8 image differs by more than one grey level"
        assert (lv > 0).mean() < 0.02, f"{(lv > 0).mean():.3%} of pixels differ by one grey level (rounding boundary)"


def test_reference_api_numpy_roundtrip():
    from music_style_transfer_ldm_b200.audio_processor import AudioPreprocessor
    ap = AudioPreprocessor()
    y = _waves(1, 66150, seed=4)[0].numpy()
    db = ap.get_mel_spectogram(y, 22050, n_mels=128)           # reference call shape (build_dataset.py:54)
    assert isinstance(db, np.ndarray) and db.shape == (128, 130)
    assert db.max() == pytest.approx(0.0, abs=1e-4) and db.min() >= -80.0
    im = ap.mel_spectogram_to_grayscale_image(db)
    u8 = np.asarray(im)
    assert u8.dtype == np.uint8 and u8.shape == (128, 130)
    ref = MO.to_grayscale_u8(MO.get_mel_spectogram(y, 22050, 128))
    assert np.abs(u8.astype(int) - ref.astype(int)).max() <= 1


def test_silence_and_errors():
    from music_style_transfer_ldm_b200.audio_processor import AudioPreprocessor
    from music_style_transfer_ldm_b200.ops import MstkError
    ap = AudioPreprocessor()
    img = ap.spectrogram_batch(torch.zeros(2, 66150).cuda(), n_mels=128, crop_frames=128)
    assert torch.all(img == 1.0)              # every bin at the reference level -> 0 dB -> 255
    with pytest.raises(MstkError):
        ap.spectrogram_batch(torch.zeros(2, 66150).cuda(), n_mels=128, crop_frames=500)
