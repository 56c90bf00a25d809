"""-m gpu: the mel front end (csrc/mel.cu) and the inverse path (csrc/griffinlim.cu) on ragged clip lengths, frame counts, band
counts and batch sizes, against the numpy oracles (oracle/mel_oracle.py, oracle/audio_inverse_oracle.py -- both restate librosa,
parity unpinned, see their headers).  The fixed tests use 3 s / 10 s clips and 128 / 130 frames; here: lengths on both sides of
the frame and hop boundaries (the last frame's zero padding, one-frame clips), prime batch sizes, 40 to 256 mel bands."""
import numpy as np
import pytest
import torch

from oracle import mel_oracle as MO
from test_gpu_mel import _waves

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("b,n,n_mels", [(1, 1025, 128), (2, 2047, 128), (3, 2048, 64), (1, 2049, 128), (5, 4097, 40), (2, 30001, 256),
                                         (7, 12345, 128), (1, 77777, 80), (3, 512 * 37, 128), (2, 512 * 37 - 1, 128)])
def test_mel_frontend_ragged_lengths(b, n, n_mels):
    from music_style_transfer_ldm_b200.audio_processor import AudioPreprocessor
    ap = AudioPreprocessor()
    w = _waves(b, n, seed=n % 97)
    db = ap.get_mel_spectogram(w.cuda(), 22050, n_mels=n_mels).cpu().numpy()
    frames = 1 + n // 512
    assert db.shape == (b, n_mels, frames)
    crop = min(frames, 384)
    img = ap.spectrogram_batch(w.cuda(), n_mels=n_mels, crop_frames=crop).cpu().numpy()
    assert img.shape == (b, 1, n_mels, crop)
    for i in range(b):
        ref_db = MO.get_mel_spectogram(w[i].numpy(), 22050, n_mels)
        np.testing.assert_allclose(db[i], ref_db, atol=2e-2, err_msg=f"clip {i} of {b} x {n}, {n_mels} bands")
        lv = np.abs(np.rint(img[i, 0] * 255) - np.rint(MO.to_unit_tensor(MO.to_grayscale_u8(ref_db), crop) * 255))
        assert lv.max() <= 1 and (lv > 0).mean() < 0.03, f"grey levels: max {lv.max()}, {(lv > 0).mean():.3%} differ"


@pytest.mark.parametrize("b,frames,n_iter", [(1, 2, 2), (2, 3, 2), (3, 17, 3), (1, 129, 2), (5, 64, 2), (2, 257, 1)])
def test_griffinlim_ragged_frame_counts(b, frames, n_iter):
    """mstk_griffinlim against the oracle's griffinlim on the STFT magnitudes of synthetic clips, few iterations (the phase
    projection a / |a| amplifies fp32-FFT rounding over the recurrence, see DESIGN section 6: 1e-3 of the waveform's norm)"""
    from oracle import audio_inverse_oracle as AO
    from music_style_transfer_ldm_b200 import ops
    w = _waves(b, 512 * (frames - 1), seed=frames).numpy().astype(np.float64)
    S = np.stack([np.abs(AO.stft(w[i])) for i in range(b)])                           # [b, 1025, frames]
    assert S.shape == (b, 1025, frames)
    U = np.random.default_rng(frames).random(S.shape).astype(np.float32)
    mag = torch.from_numpy(np.ascontiguousarray(S.transpose(0, 2, 1)).astype(np.float32)).cuda()
    u = torch.from_numpy(np.ascontiguousarray(U.transpose(0, 2, 1))).cuda()
    out = ops.griffinlim(mag, u, n_iter=n_iter, momentum=0.99).cpu().numpy()
    assert out.shape == (b, 512 * (frames - 1))
    for i in range(b):
        ref = AO.griffinlim(S[i].astype(np.float32).astype(np.float64), U[i].astype(np.float64), n_iter=n_iter, momentum=0.99)
        assert ref.shape == out[i].shape
        err = np.linalg.norm(out[i] - ref) / max(np.linalg.norm(ref), 1e-30)
        assert err <= 1e-3, f"clip {i} of {b} x {frames} frames after {n_iter} iterations: {err:.2e}"
