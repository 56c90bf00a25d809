"""Drop-in surface of the reference's data/audio_processor.py forward path on libmstk.

    AudioPreprocessor(target_sr=22050)
        .get_mel_spectogram(audio, sr, n_mels=256)            audio_processor.py:44-53
        .mel_spectogram_to_grayscale_image(spectogram, 80)    audio_processor.py:55-73
        .spectrogram_batch(waves, n_mels, crop_frames)        (batched device path: the two calls above
                                                               + dataset.py:51-53 crop and /255 in one go)
        .grayscale_mel_spectogram_image_to_audio(image, sr, im_height, im_width, max_db=80)     audio_processor.py:81-100
        .grayscale_spectogram_image_to_audio(image, im_height, im_width, max_db=80, n_iter=32)  audio_processor.py:137-162
        .audio_batch(images, sr, max_db, n_iter)              (batched device path of the inverse: images -> waveforms)

The arithmetic (librosa.feature.melspectrogram + power_to_db(ref=np.max) + uint8 image
quantisation) runs in the mel kernels of csrc/mel.cu.  The only host work is building the
Slaney mel filterbank, a constant of (sr, n_mels), once.  mp3 decoding / resampling
(load_audio) and trim_silence are outside the hot path and are not provided.
The inverse (dB image -> mel power -> NNLS magnitude -> 32 Griffin-Lim iterations)
runs in csrc/griffinlim.cu; its host work is the pseudo-inverse of the filterbank, once per (sr, n_mels).
There is no CPU path: waveforms and images are moved to the CUDA device.
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
        self._pinv: Dict[Tuple[int, int, int], Tuple[torch.Tensor, float]] = {}

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

    # -- inverse path (audio_processor.py:81-100, 137-162) ---------------------------------------
    PGTOL = 1e-5          # scipy.optimize.fmin_l_bfgs_b's default, which librosa.util.nnls leaves in place

    def _pseudo_inverse(self, sr: int, n_mels: int, device) -> Tuple[torch.Tensor, float]:
        key = (int(sr), int(n_mels), torch.device(device).index or 0)
        ent = self._pinv.get(key)
        if ent is None:
            fb = slaney_filterbank(int(sr), int(n_mels))
            pinv = np.linalg.pinv(fb).astype(np.float32)            # librosa.util.nnls: np.linalg.pinv(A), A float32
            lip = float(np.linalg.norm(fb.astype(np.float64), 2) ** 2)
            ent = (torch.from_numpy(pinv).to(device).contiguous(), 1.0 / lip)
            self._pinv[key] = ent
        return ent

    @staticmethod
    def _image_u8(image, im_height, im_width) -> np.ndarray:
        """audio_processor.py:93-95: the image's raw bytes as [im_height, im_width] uint8"""
        raw = image.tobytes() if hasattr(image, "tobytes") else bytes(image)
        return np.frombuffer(raw, dtype=np.uint8).reshape(im_height, im_width)

    def mel_images_to_magnitude(self, images_u8: torch.Tensor, sr=None, max_db=80) -> torch.Tensor:
        """uint8 [b, n_mels, frames] on the device -> STFT magnitude [b, frames, 1025] (librosa.feature.inverse.mel_to_stft).
        librosa's NNLS is L-BFGS-B started at clip(pinv(A) B, 0); for a dB image the projected-gradient test passes at
        that point and L-BFGS-B returns it unchanged, which is what the kernel computes.  The test is evaluated on the
        device for every clip; should it fail (an input outside the dB-image domain) projected-gradient steps are run
        until it passes -- not the reference's iterates, and said so here."""
        n_mels = images_u8.shape[1]
        fb, ranges = self._filterbank(sr or self.target_sr, n_mels, images_u8.device)
        pinv, step = self._pseudo_inverse(sr or self.target_sr, n_mels, images_u8.device)
        mag, pg = ops.mel_to_stft(images_u8, pinv, fb, ranges, max_db=max_db)
        iters = 64
        while float(pg.max()) > self.PGTOL and iters <= 4096:
            mag, pg = ops.mel_to_stft(images_u8, pinv, fb, ranges, max_db=max_db, pg_iters=iters, pg_step=step)
            iters *= 4
        return mag

    def audio_batch(self, images, sr=None, max_db=80, n_iter=32, momentum=0.99, phase_u=None) -> torch.Tensor:
        """Batched device path of the inverse: images [b, 1, n_mels, frames] / [b, n_mels, frames], uint8 or fp32 in [0,1]
        (quantised as the forward path does, trunc(255 x + 0.5)) -> waveforms [b, 512 * (frames - 1)] on the device.
        ``phase_u`` [b, frames, 1025] in [0,1) fixes the initial phases 2 pi u (default: torch.rand on the device; the
        reference draws them from numpy's default_rng)."""
        if not torch.cuda.is_available():
            raise MstkError("AudioPreprocessor: no CUDA device; the inverse audio path runs on sm_100a kernels only")
        x = images if isinstance(images, torch.Tensor) else torch.as_tensor(np.asarray(images))
        if x.dim() == 4 and x.shape[1] == 1:
            x = x[:, 0]
        if x.dim() != 3:
            raise MstkError(f"audio_batch: expected [b, n_mels, frames] or [b, 1, n_mels, frames], got {tuple(x.shape)}")
        x = x.to("cuda", non_blocking=True) if not x.is_cuda else x
        if x.dtype != torch.uint8:
            x = (x.detach().float().clamp(0, 1) * 255.0 + 0.5).to(torch.uint8)
        x = x.contiguous()
        mag = self.mel_images_to_magnitude(x, sr=sr, max_db=max_db)
        if phase_u is None:
            phase_u = torch.rand(mag.shape, dtype=torch.float32, device=mag.device)
        return ops.griffinlim(mag, phase_u.to(mag.device, torch.float32).contiguous(), n_iter=n_iter, momentum=momentum)

    def grayscale_mel_spectogram_image_to_audio(self, image, sr, im_height, im_width, max_db=80, phase_u=None):
        """audio_processor.py:81-100: grayscale image of a mel spectrogram -> waveform (numpy float32, as the reference)."""
        u8 = torch.from_numpy(self._image_u8(image, im_height, im_width).copy())[None]
        return self.audio_batch(u8, sr=sr, max_db=max_db, phase_u=phase_u)[0].cpu().numpy()

    def grayscale_spectogram_image_to_audio(self, image, im_height, im_width, max_db=80, n_iter=32, hop_length=None, win_length=None,
                                            phase_u=None):
        """audio_processor.py:137-162: grayscale image of a dB STFT magnitude -> waveform by Griffin-Lim.  The kernels are
        built for librosa's defaults at im_height = 1025 (n_fft 2048, hop 512, full-length Hann); anything else raises."""
        if im_height != 1025 or hop_length not in (None, HOP_LENGTH) or win_length not in (None, N_FFT):
            raise MstkError("grayscale_spectogram_image_to_audio: only im_height 1025 (n_fft 2048), hop_length 512, win_length 2048")
        if not torch.cuda.is_available():
            raise MstkError("AudioPreprocessor: no CUDA device; the inverse audio path runs on sm_100a kernels only")
        u8 = torch.from_numpy(self._image_u8(image, im_height, im_width).copy())[None].to("cuda")
        mag = ops.db_image_to_mag(u8, max_db=max_db)
        if phase_u is None:
            phase_u = torch.rand(mag.shape, dtype=torch.float32, device=mag.device)
        return ops.griffinlim(mag, phase_u.to(mag.device, torch.float32).contiguous(), n_iter=n_iter)[0].cpu().numpy()
