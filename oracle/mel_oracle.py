"""numpy CPU restatement of the reference's mel front end.  TEST INFRASTRUCTURE ONLY.

Follows data/audio_processor.py:44-73 of the reference:

    mel_spec     = librosa.feature.melspectrogram(y=audio, sr=sr, n_mels=n_mels)   # :51
    log_mel_spec = librosa.power_to_db(mel_spec, ref=np.max)                       # :52
    image        = uint8(clip((S + 80) * 255/80, 0, 255) + 0.5)                    # :63-69
    tensor       = crop[:128, :128] / 255                                          # models/dataset.py:51-53

PARITY UNPINNED: the arithmetic lives in librosa==0.10.2.post1 (requirements.txt:40),
which is neither vendored under /root/reference nor installed, and the reference holds no
golden vector for it (tests/audio_reconstruction_test.py has no asserts).  This file restates
librosa's published algorithm and defaults (n_fft=2048, hop_length=512, win_length=n_fft,
window='hann' (periodic), center=True, pad_mode='constant', power=2.0; filters.mel with
fmin=0, fmax=sr/2, htk=False, norm='slaney', float32; power_to_db amin=1e-10, top_db=80)
and is cross-checked against torchaudio's independent implementation in
tests/test_mel_oracle.py.
"""
from __future__ import annotations

import numpy as np

N_FFT = 2048
HOP = 512


def hz_to_mel(f):
    """librosa.core.convert.hz_to_mel(htk=False): Slaney's auditory-toolbox scale."""
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep, mels)


def mel_to_hz(m):
    """librosa.core.convert.mel_to_hz(htk=False)."""
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_filterbank(sr: int = 22050, n_fft: int = N_FFT, n_mels: int = 128) -> np.ndarray:
    """librosa.filters.mel(sr, n_fft, n_mels, fmin=0, fmax=sr/2, htk=False, norm='slaney') -> [n_mels, 1+n_fft/2] fp32."""
    fmax = sr / 2.0
    fftfreqs = np.linspace(0.0, sr / 2.0, 1 + n_fft // 2)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(0.0), hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    w = np.zeros((n_mels, 1 + n_fft // 2), dtype=np.float64)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0.0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    w *= enorm[:, None]
    return w.astype(np.float32)


def hann_periodic(n: int = N_FFT) -> np.ndarray:
    """scipy.signal.get_window('hann', n, fftbins=True)"""
    return (0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)).astype(np.float32)


def stft_power(y: np.ndarray, n_fft: int = N_FFT, hop: int = HOP) -> np.ndarray:
    """|librosa.stft(y, n_fft, hop, center=True, pad_mode='constant')|**2 -> [1+n_fft/2, frames] fp32."""
    y = np.asarray(y, dtype=np.float32)
    yp = np.pad(y, (n_fft // 2, n_fft // 2), mode="constant")
    frames = 1 + (len(yp) - n_fft) // hop
    idx = np.arange(n_fft)[None, :] + hop * np.arange(frames)[:, None]
    fr = yp[idx] * hann_periodic(n_fft)[None, :]
    spec = np.fft.rfft(fr.astype(np.float32), axis=1).astype(np.complex64)
    return (np.abs(spec) ** 2).T.astype(np.float32)


def melspectrogram(y: np.ndarray, sr: int = 22050, n_mels: int = 128) -> np.ndarray:
    """librosa.feature.melspectrogram(y=y, sr=sr, n_mels=n_mels) -> [n_mels, frames] fp32."""
    return (mel_filterbank(sr, N_FFT, n_mels) @ stft_power(y)).astype(np.float32)


def power_to_db(S: np.ndarray, amin: float = 1e-10, top_db: float = 80.0) -> np.ndarray:
    """librosa.power_to_db(S, ref=np.max)."""
    S = np.asarray(S)
    ref = np.max(S)
    log_spec = 10.0 * np.log10(np.maximum(amin, S))
    log_spec -= 10.0 * np.log10(np.maximum(amin, ref))
    return np.maximum(log_spec, log_spec.max() - top_db)


def get_mel_spectogram(audio: np.ndarray, sr: int = 22050, n_mels: int = 128) -> np.ndarray:
    """AudioPreprocessor.get_mel_spectogram (audio_processor.py:44-53)."""
    return power_to_db(melspectrogram(audio, sr, n_mels))


def to_grayscale_u8(spectogram: np.ndarray, max_db: float = 80) -> np.ndarray:
    """AudioPreprocessor.mel_spectogram_to_grayscale_image (audio_processor.py:55-73), as the uint8 array."""
    s = spectogram + max_db
    s = s * (255.0 / max_db)
    s = np.clip(s, 0, 255)
    return (s + 0.5).astype(np.uint8)


def to_unit_tensor(u8: np.ndarray, crop: int) -> np.ndarray:
    """models/dataset.py:51-53: top-left crop + ToTensor's /255."""
    return u8[:, :crop].astype(np.float32) / 255.0


def waveform_to_unit_image(audio: np.ndarray, sr: int = 22050, n_mels: int = 128, crop_frames: int = 128) -> np.ndarray:
    """whole front end: waveform -> [n_mels, crop_frames] fp32 in [0,1]"""
    return to_unit_tensor(to_grayscale_u8(get_mel_spectogram(audio, sr, n_mels)), crop_frames)
