"""CPU: the inverse-audio oracle (oracle/audio_inverse_oracle.py, a restatement of librosa's mel_to_audio / griffinlim;
PARITY UNPINNED, librosa is absent) against independent implementations that ARE here: torch.stft / torch.istft for the
transforms, scipy's L-BFGS-B for the claim that librosa's NNLS returns its starting point on a dB image."""
import numpy as np
import pytest
import torch

from oracle import audio_inverse_oracle as AO
from oracle import mel_oracle as MO


def _image(frames=40, seed=0):
    rng = np.random.default_rng(seed)
    n = 512 * (frames - 1)
    t = np.arange(n) / 22050.0
    y = (0.3 * np.sin(2 * np.pi * 440 * t) + 0.2 * np.sin(2 * np.pi * 1320 * t * (1 + 0.1 * t)) + 0.05 * rng.standard_normal(n)).astype(np.float32)
    return MO.to_grayscale_u8(MO.get_mel_spectogram(y, 22050, 128)), y


def test_stft_istft_match_torch():
    rng = np.random.default_rng(3)
    y = rng.standard_normal(512 * 19).astype(np.float32)
    win = torch.hann_window(2048, periodic=True)
    ref = torch.stft(torch.from_numpy(y), 2048, 512, 2048, win, center=True, pad_mode="constant", return_complex=True).numpy()
    got = AO.stft(y)
    assert got.shape == ref.shape == (1025, 20)
    assert np.abs(got - ref).max() <= 2e-4 * np.abs(ref).max()
    spec = (rng.standard_normal((1025, 20)) + 1j * rng.standard_normal((1025, 20))).astype(np.complex64)
    spec[0].imag = 0; spec[-1].imag = 0
    ref_y = torch.istft(torch.from_numpy(spec), 2048, 512, 2048, win, center=True, length=512 * 19).numpy()
    got_y = AO.istft(spec)
    assert got_y.shape == ref_y.shape
    assert np.abs(got_y - ref_y).max() <= 1e-5 * np.abs(ref_y).max() + 1e-7
    # perfect reconstruction away from the trimmed edges
    back = AO.istft(AO.stft(y))
    assert np.abs(back - y)[1024:-1024].max() <= 1e-5


def test_nnls_returns_its_starting_point_on_a_db_image():
    pytest.importorskip("scipy")
    img, _ = _image(24)
    M = AO.db_to_power(AO.image_to_db(img))
    assert M.max() <= 1.0
    A = MO.mel_filterbank(22050, 2048, 128)
    x0 = AO.nnls(A, M, use_lbfgs=False)
    assert AO.projected_gradient_norm(x0, A, M) < 1e-5          # scipy's pgtol: L-BFGS-B stops before its first iteration
    x1 = AO.nnls(A, M, use_lbfgs=True)
    assert np.array_equal(x0, x1)
    assert x0.min() >= 0 and x0.shape == (1025, 24)


def test_griffinlim_converges_and_is_deterministic_in_the_phases():
    img, _ = _image(24, seed=1)
    S = AO.mel_to_stft(AO.db_to_power(AO.image_to_db(img)), use_lbfgs=False)
    U = np.random.default_rng(5).random(S.shape)
    a0 = AO.griffinlim(S, U, n_iter=0)
    a8 = AO.griffinlim(S, U, n_iter=8)
    a32 = AO.griffinlim(S, U, n_iter=32)
    assert a32.shape == (512 * 23,) and a32.dtype == np.float32
    c0, c8, c32 = (AO.spectral_convergence(a, S) for a in (a0, a8, a32))
    assert c32 < c8 < c0 and c32 < 0.5
    assert np.array_equal(a32, AO.griffinlim(S, U, n_iter=32))
    # the whole reference call
    full = AO.grayscale_mel_spectogram_image_to_audio(img, U, use_lbfgs=False)
    assert np.array_equal(full, a32)


def test_window_sumsquare_interior_is_three_halves():
    w = AO.window_sumsquare(12)
    assert w.shape == (2048 + 512 * 11,)
    assert np.abs(w[1536:-1536] - 1.5).max() < 1e-6
    assert w[0] == 0.0 and w[1] > 0
