"""Mel front end, CPU side: the numpy restatement of librosa's algorithm (oracle/mel_oracle.py,
PARITY UNPINNED -- librosa is absent and the reference holds no golden vector) cross-checked
against torchaudio's independent implementation, and the product's host-side filterbank
builder against both."""
import numpy as np
import pytest
import torch

from oracle import mel_oracle as MO


def _wave(seed, n):
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(n) / 22050.0
    # a few partials + noise: realistic dynamic range (pure noise never reaches the -80 dB floor)
    y = 0.3 * torch.sin(2 * np.pi * 220 * t) + 0.1 * torch.sin(2 * np.pi * 1760 * t) * torch.exp(-3 * t)
    return (y + 0.01 * torch.randn(n, generator=g)).numpy().astype(np.float32)


def test_frame_counts():
    # 3 s -> 130 frames, 10 s -> 431 frames
    assert MO.stft_power(np.zeros(66150, np.float32)).shape == (1025, 130)
    assert MO.stft_power(np.zeros(220500, np.float32)).shape == (1025, 431)


def test_filterbank_vs_torchaudio():
    ta = pytest.importorskip("torchaudio")
    fb = ta.functional.melscale_fbanks(1025, 0.0, 11025.0, 128, 22050, norm="slaney", mel_scale="slaney").T.numpy()
    np.testing.assert_allclose(MO.mel_filterbank(22050, 2048, 128), fb, atol=2e-7, rtol=1e-4)


def test_product_filterbank_matches_oracle():
    from music_style_transfer_ldm_b200.audio_processor import slaney_filterbank
    for n_mels in (128, 256):
        np.testing.assert_allclose(slaney_filterbank(22050, n_mels), MO.mel_filterbank(22050, 2048, n_mels), atol=1e-7, rtol=1e-5)


def test_melspectrogram_vs_torchaudio():
    ta = pytest.importorskip("torchaudio")
    y = _wave(3, 66150)
    tr = ta.transforms.MelSpectrogram(sample_rate=22050, n_fft=2048, hop_length=512, n_mels=128, center=True,
                                      pad_mode="constant", power=2.0, norm="slaney", mel_scale="slaney")
    ref = tr(torch.from_numpy(y)).numpy()
    got = MO.melspectrogram(y, 22050, 128)
    assert got.shape == ref.shape == (128, 130)
    np.testing.assert_allclose(got, ref, rtol=2e-4, atol=1e-7 * ref.max())


def test_db_and_image_properties():
    y = _wave(5, 66150)
    db = MO.get_mel_spectogram(y, 22050, 128)
    assert db.max() == pytest.approx(0.0, abs=1e-5) and db.min() >= -80.0
    u8 = MO.to_grayscale_u8(db)
    assert u8.dtype == np.uint8 and u8.max() == 255
    img = MO.waveform_to_unit_image(y, crop_frames=128)
    assert img.shape == (128, 128) and 0.0 <= img.min() and img.max() == 1.0
    # silence: every bin is at the reference level -> 0 dB -> 255
    assert (MO.to_grayscale_u8(MO.get_mel_spectogram(np.zeros(4096, np.float32))) == 255).all()
