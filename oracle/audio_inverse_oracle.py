"""numpy / scipy CPU restatement of the reference's inverse audio path.  TEST INFRASTRUCTURE ONLY.

Follows data/audio_processor.py:81-100 of the reference (and :137-162 for the plain-STFT variant):

    image_bytes  = frombuffer(image.tobytes(), uint8).reshape(im_height, im_width)          # :93-95
    log_mel_spec = image_bytes.astype(float32) * (max_db / 255.0) - max_db                   # :96
    mel_spec     = librosa.db_to_power(log_mel_spec)                                         # :97
    audio        = librosa.feature.inverse.mel_to_audio(mel_spec, sr=sr)                     # :98

PARITY UNPINNED: the arithmetic lives in librosa==0.10.2.post1 (requirements.txt:40), which is neither
vendored under /root/reference nor installed, and the reference holds no golden vector for it
(tests/audio_reconstruction_test.py has no asserts).  This file restates librosa's published algorithm
and defaults:

  * feature.inverse.mel_to_stft: mel_basis = filters.mel(sr, n_fft=2048, n_mels=M.shape[-2], dtype=M.dtype);
    inverse = util.nnls(mel_basis, M); return inverse ** (1 / power), power = 2.
  * util.nnls (2-D B, columns <= MAX_MEM_BLOCK // (rows * itemsize) = 512): _nnls_lbfgs_block:
    x_init = clip(pinv(A) @ B, 0, None); scipy.optimize.fmin_l_bfgs_b(_nnls_obj, x_init, bounds = [(0, None)] * size,
    m = A.shape[1]) with _nnls_obj = 0.5 * sum((A x - B)^2) / B.size and its gradient A^T (A x - B) / B.size.
    scipy's defaults (factr 1e7, pgtol 1e-5) apply: for a dB image (power <= 1) the projected gradient at x_init is
    already below pgtol (|A| <= 0.039, |A x - B| <= ~1, B.size = 16384 -> <= 5e-6), so L-BFGS-B returns x_init.
  * griffinlim(S, n_iter=32, hop_length=n_fft // 4, window='hann', center=True, pad_mode='constant', momentum=0.99,
    init='random'): angles = exp(2 pi i U) * S; per iteration inverse = istft(angles); rebuilt = stft(inverse);
    angles = rebuilt - momentum / (1 + momentum) * tprev (from the second iteration on); angles /= |angles| + tiny;
    angles *= S; tprev = rebuilt.  Returns istft(angles).  U ~ default_rng().random(S.shape): the caller passes U so
    that two implementations can be compared on the same phases.
  * istft: irfft of every frame * hann, overlap-add at hop 512, division by the window sum-square where it exceeds
    tiny(float32), trimmed by n_fft // 2 at both ends -> hop * (frames - 1) samples.
"""
from __future__ import annotations

import numpy as np

from .mel_oracle import N_FFT, HOP, hann_periodic, mel_filterbank

TINY32 = float(np.finfo(np.float32).tiny)


def image_to_db(image_u8: np.ndarray, max_db: float = 80) -> np.ndarray:
    """audio_processor.py:93-96 (and :150-153): uint8 image -> dB, float32."""
    return np.asarray(image_u8, dtype=np.uint8).astype(np.float32) * (max_db / 255.0) - max_db


def db_to_power(s_db: np.ndarray) -> np.ndarray:
    """librosa.db_to_power(S_db, ref=1.0) = ref * 10 ** (0.1 * S_db)"""
    return np.power(10.0, 0.1 * s_db).astype(np.float32)


def db_to_amplitude(s_db: np.ndarray) -> np.ndarray:
    """librosa.db_to_amplitude(S_db, ref=1.0) = db_to_power(S_db, ref ** 2) ** 0.5"""
    return (db_to_power(s_db) ** 0.5).astype(np.float32)


def mel_pinv(sr: int = 22050, n_mels: int = 128) -> np.ndarray:
    """np.linalg.pinv(filters.mel(..., dtype=float32)) -> [1 + n_fft/2, n_mels] float32 (as util.nnls computes it)."""
    return np.linalg.pinv(mel_filterbank(sr, N_FFT, n_mels)).astype(np.float32)


def nnls_obj(x: np.ndarray, shape, A: np.ndarray, B: np.ndarray):
    """librosa.util._nnls._nnls_obj"""
    x = x.reshape(shape)
    diff = np.einsum("mf,...ft->...mt", A, x, optimize=True) - B
    value = (1 / B.size) * 0.5 * np.sum(diff ** 2)
    grad = (1 / B.size) * np.einsum("mf,...mt->...ft", A, diff, optimize=True)
    return value, grad.flatten()


def projected_gradient_norm(x: np.ndarray, A: np.ndarray, B: np.ndarray) -> float:
    """L-BFGS-B's `projgr` for the bounds (0, None): the quantity compared with pgtol before the first iteration."""
    _, g = nnls_obj(x.flatten(), x.shape, A, B)
    g = g.reshape(x.shape)
    return float(np.max(np.abs(np.where(g >= 0, np.minimum(x, g), g))))


def nnls(A: np.ndarray, B: np.ndarray, use_lbfgs: bool = True) -> np.ndarray:
    """librosa.util.nnls for 2-D B of at most 512 columns (one block)."""
    import scipy.optimize
    assert B.ndim == 2 and B.shape[-1] <= (2 ** 18) // (B.shape[0] * A.itemsize)
    x_init = np.einsum("fm,...mt->...ft", np.linalg.pinv(A), B, optimize=True)
    np.clip(x_init, 0, None, out=x_init)
    if not use_lbfgs:
        return x_init.astype(A.dtype)
    shape = x_init.shape
    x, _, _ = scipy.optimize.fmin_l_bfgs_b(nnls_obj, x_init, args=(shape, A, B), bounds=[(0, None)] * x_init.size, m=A.shape[1])
    return x.reshape(shape).astype(A.dtype)


def mel_to_stft(M: np.ndarray, sr: int = 22050, power: float = 2.0, use_lbfgs: bool = True) -> np.ndarray:
    """librosa.feature.inverse.mel_to_stft -> [1025, frames] float32 magnitudes."""
    M = np.asarray(M, dtype=np.float32)
    basis = mel_filterbank(sr, N_FFT, M.shape[-2])
    inverse = nnls(basis, M, use_lbfgs)
    return np.power(inverse, 1.0 / power, out=inverse)


def window_sumsquare(n_frames: int) -> np.ndarray:
    """librosa.filters.window_sumsquare('hann', n_frames, hop 512, n_fft 2048, norm=None, dtype=float32)"""
    n = N_FFT + HOP * (n_frames - 1)
    x = np.zeros(n, dtype=np.float32)
    win_sq = (hann_periodic(N_FFT).astype(np.float64) ** 2)
    for i in range(n_frames):
        s = i * HOP
        x[s:min(n, s + N_FFT)] += win_sq[:max(0, min(N_FFT, n - s))]          # float32 array += float64 terms
    return x


def istft(spec: np.ndarray) -> np.ndarray:
    """librosa.istft(stft_matrix [1025, frames] complex64, hop 512, window hann, center=True) -> float32 [hop * (frames-1)]"""
    n_frames = spec.shape[1]
    win = hann_periodic(N_FFT)
    ytmp = win[:, None] * np.fft.irfft(spec, n=N_FFT, axis=0).astype(np.float32)
    n = N_FFT + HOP * (n_frames - 1)
    y = np.zeros(n, dtype=np.float32)
    for f in range(n_frames):
        y[f * HOP:f * HOP + N_FFT] += ytmp[:, f]
    wss = window_sumsquare(n_frames)
    nz = wss > TINY32
    y[nz] /= wss[nz]
    return y[N_FFT // 2: N_FFT // 2 + HOP * (n_frames - 1)]


def stft(y: np.ndarray) -> np.ndarray:
    """librosa.stft(y, n_fft 2048, hop 512, hann, center=True, pad_mode='constant') -> complex64 [1025, frames]"""
    yp = np.pad(np.asarray(y, dtype=np.float32), (N_FFT // 2, N_FFT // 2), mode="constant")
    frames = 1 + (len(yp) - N_FFT) // HOP
    idx = np.arange(N_FFT)[None, :] + HOP * np.arange(frames)[:, None]
    fr = yp[idx] * hann_periodic(N_FFT)[None, :]
    return np.fft.rfft(fr.astype(np.float32), axis=1).astype(np.complex64).T


def griffinlim(S: np.ndarray, U: np.ndarray, n_iter: int = 32, momentum: float = 0.99) -> np.ndarray:
    """librosa.griffinlim(S, n_iter=32, momentum=0.99, init='random') with the uniform draws U in [0,1) given ([1025, frames])."""
    S = np.asarray(S, dtype=np.float32)
    angles = np.exp(2j * np.pi * np.asarray(U, dtype=np.float64)).astype(np.complex64)     # util.phasor(2 pi U)
    angles *= S
    tprev = None
    for _ in range(n_iter):
        inverse = istft(angles)
        rebuilt = stft(inverse)
        angles = rebuilt.copy()
        if tprev is not None:
            angles -= np.complex64(momentum / (1 + momentum)) * tprev
        angles /= np.abs(angles) + TINY32
        angles *= S
        tprev = rebuilt
    return istft(angles)


def grayscale_mel_spectogram_image_to_audio(image_u8: np.ndarray, U: np.ndarray, sr: int = 22050, max_db: float = 80,
                                            n_iter: int = 32, use_lbfgs: bool = True) -> np.ndarray:
    """AudioPreprocessor.grayscale_mel_spectogram_image_to_audio (audio_processor.py:81-100); image [n_mels, frames] uint8."""
    mel_spec = db_to_power(image_to_db(image_u8, max_db))
    S = mel_to_stft(mel_spec, sr=sr, use_lbfgs=use_lbfgs)
    return griffinlim(S, U, n_iter=n_iter)


def grayscale_spectogram_image_to_audio(image_u8: np.ndarray, U: np.ndarray, max_db: float = 80, n_iter: int = 32) -> np.ndarray:
    """AudioPreprocessor.grayscale_spectogram_image_to_audio (audio_processor.py:137-162); image [1025, frames] uint8."""
    return griffinlim(db_to_amplitude(image_to_db(image_u8, max_db)), U, n_iter=n_iter)


def spectral_convergence(audio: np.ndarray, S: np.ndarray) -> float:
    """|| |STFT(audio)| - S ||_F / || S ||_F: the quantity Griffin-Lim drives down (implementation-independent check)."""
    R = np.abs(stft(audio))
    return float(np.linalg.norm(R - S) / np.linalg.norm(S))
