"""Drop-in surface of the reference's data/audio_processor.py forward path on libmstk.

    AudioPreprocessor(target_sr=22050)
        .get_mel_spectogram(audio, sr, n_mels=256)            audio_processor.py:44-53
        .mel_spectogram_to_grayscale_image(spectogram, 80)    audio_processor.py:55-73
        .spectrogram_batch(waves, n_mels, crop_frames)        (batched device path: the two calls above
                                                               + dataset.py:51-53 crop and /255 in one go)

The arithmetic (librosa.feature.melspectrogram + power_to_db(ref=np.max) + uint8 image
quantisation) runs in the mel kernels of csrc/mel.cu.  The only host work is building the
Slaney mel filterbank, a constant of (sr, n_mels), once.  mp3 decoding / resampling
(load_audio), trim_silence and the Griffin-Lim inverse are outside the hot path (SURVEY section 2
row 5) and are not provided.  There is no CPU path: waveforms are moved to the CUDA device.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np
import torch

from . import ops
from .ops import MstkError

N_FFT = 2048
HOP_LENGTH = 512


def _slaney_hz_to_mel(f: np.ndarray) -> np.ndarray:
    f = np.asarray(f, dtype=np.float64)
    lin = f * (3.0 / 200.0)
    with np.errstate(divide="ignore"):
        log = 15.0 + np.log(np.where(f > 0, f, 1.0) / 1000.0) * (27.0 / math.log(6.4))
    return np.where(f >= 1000.0, log, lin)


def _slaney_mel_to_hz(m: np.ndarray) -> np.ndarray:
    m = np.asarray(m, dtype=np.float64)
    return np.where(m >= 15.0, 1000.0 * np.exp((m - 15.0) * (math.log(6.4) / 27.0)), m * (200.0 / 3.0))


def slaney_filterbank(sr: int, n_mels: int, n_fft: int = N_FFT) -> np.ndarray:
    """Triangular, area-normalised ('slaney') mel filters on the Slaney scale, fmin 0, fmax sr/2:
    what librosa.filters.mel returns for the reference's call.  [n_mels, n_fft/2+1] fp32."""
    nb = n_fft // 2 + 1
    bins = np.arange(nb, dtype=np.float64) * (sr / 2.0) / (nb - 1)
    edges = _slaney_mel_to_hz(np.linspace(0.0, float(_slaney_hz_to_mel(np.array(sr / 2.0))), n_mels + 2))
    lo, ce, hi = edges[:-2, None], edges[1:-1, None], edges[2:, None]
    up = (bins[None, :] - lo) / (ce - lo)
    down = (hi - bins[None, :]) / (hi - ce)
    tri = np.clip(np.minimum(up, down), 0.0, None)
    return (tri * (2.0 / (hi - lo))).astype(np.float32)


class AudioPreprocessor:
    def __init__(self, target_sr=22050):
        self.target_sr = target_sr
        self._fb: Dict[Tuple[int, int, int], Tuple[torch.Tensor, torch.Tensor]] = {}

    # -- helpers ---------------------------------------------------------------------------
    def _filterbank(self, sr: int, n_mels: int, device) -> torch.Tensor:
        key = (int(sr), int(n_mels), torch.device(device).index or 0)
        fb = self._fb.get(key)
        if fb is None:
            fbm = torch.from_numpy(slaney_filterbank(int(sr), int(n_mels))).to(device).contiguous()
            fb = (fbm, ops.mel_band_ranges(fbm))          # the non-zero bin range of every band, once per filterbank
            self._fb[key] = fb
        return fb

    @staticmethod
    def _device_waves(audio) -> Tuple[torch.Tensor, bool]:
        if not torch.cuda.is_available():
            raise MstkError("AudioPreprocessor: no CUDA device; the mel front end runs on sm_100a kernels only")
        w = torch.as_tensor(audio) if not isinstance(audio, torch.Tensor) else audio
        single = w.dim() == 1
        if single:
            w = w[None]
        if w.dim() != 2:
            raise MstkError(f"AudioPreprocessor: expected a waveform [len] or a batch [b, len], got {tuple(w.shape)}")
        if not w.is_cuda:
            w = w.to("cuda", non_blocking=True)
        return w.detach().float().contiguous(), single

    # -- reference API -----------------------------------------------------------------------
    def get_mel_spectogram(self, audio, sr, n_mels=256):
        """Log-scaled (dB re clip max, floored at -80) mel spectrogram.  A 1-D numpy waveform
        returns a numpy [n_mels, frames] array as the reference does; CUDA tensors / batches
        return CUDA tensors [b, n_mels, frames]."""
        w, single = self._device_waves(audio)
        frames = 1 + w.shape[1] // HOP_LENGTH
        fb, ranges = self._filterbank(sr, n_mels, w.device)
        _, db = ops.mel_frontend(w, fb, crop_frames=frames, want_db=True, band_ranges=ranges)
        if isinstance(audio, torch.Tensor):
            return db[0] if single else db
        out = db.cpu().numpy()
        return out[0] if single else out

    def mel_spectogram_to_grayscale_image(self, spectogram, max_db=80):
        """uint8 image of a dB spectrogram (audio_processor.py:55-73).  Host-side convenience that
        mirrors the reference; the batched device path (`spectrogram_batch`) fuses this step
        into the mel kernels.  Returns a PIL image when PIL is importable, else the uint8 array."""
        s = spectogram.detach().cpu().numpy() if isinstance(spectogram, torch.Tensor) else np.asarray(spectogram)
        u8 = (np.clip((s + max_db) * (255.0 / max_db), 0, 255) + 0.5).astype(np.uint8)
        try:
            from PIL import Image
            return Image.fromarray(u8)
        except ImportError:       # pragma: no cover
            return u8

    # -- batched device path (config 5) ----------------------------------------------------
    def spectrogram_batch(self, waves, sr=None, n_mels=128, crop_frames=None) -> torch.Tensor:
        """waves [b, len] -> [b, 1, n_mels, crop_frames] fp32 in [0,1] on the device: mel -> dB ->
        uint8 quantisation -> top-left crop -> /255, i.e. exactly what training and sampling read
        back from the reference's PNG files (dataset.py:51-53).  ``crop_frames`` defaults to the
        largest multiple of 64 that fits (the UNet skip connections need it)."""
        w, _ = self._device_waves(waves)
        frames = 1 + w.shape[1] // HOP_LENGTH
        if crop_frames is None:
            crop_frames = (frames // 64) * 64
        if crop_frames <= 0 or crop_frames > frames:
            raise MstkError(f"spectrogram_batch: crop_frames={crop_frames} outside 1..{frames}")
        fb, ranges = self._filterbank(sr or self.target_sr, n_mels, w.device)
        out, _ = ops.mel_frontend(w, fb, crop_frames=crop_frames, band_ranges=ranges)
        return out.view(out.shape[0], 1, n_mels, crop_frames)
