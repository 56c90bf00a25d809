"""-m gpu: BASELINE config 5 in miniature -- 10 s waveforms -> mel/STFT front end (csrc/mel.cu) -> crop 128x384 ->
VAE encode -> style-transfer DDIM loop -> decode, everything on the CUDA path, against the CPU oracles
(oracle/mel_oracle.py for the front end, oracle/ldm_oracle.py for the model) on the same inputs."""
import numpy as np
import pytest
import torch

from gpu_util import assert_close_mode
from oracle import ldm_oracle as O
from oracle import mel_oracle as MO
from test_gpu_mel import _waves
from test_gpu_model import fixed_randn_like

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_waveform_to_transferred_spectrogram(sd, precision):
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200.audio_processor import AudioPreprocessor
    b, n, crop, steps = 3, 220500, 384, 6
    content_w, style_w = _waves(b, n, seed=1), _waves(b, n, seed=2).flip(0)
    ap = AudioPreprocessor()
    xc = ap.spectrogram_batch(content_w.to(DEV), n_mels=128, crop_frames=crop)        # [b,1,128,384] in [0,1]
    xs = ap.spectrogram_batch(style_w.to(DEV), n_mels=128, crop_frames=crop)
    assert xc.shape == (b, 1, 128, crop) and float(xc.min()) >= 0.0 and float(xc.max()) <= 1.0
    # front end vs the numpy restatement: at most one grey level on a few rounding-boundary pixels
    for i in range(b):
        ref = MO.to_unit_tensor(MO.to_grayscale_u8(MO.get_mel_spectogram(content_w[i].numpy(), 22050, 128)), crop)
        assert np.abs(np.rint(xc[i, 0].cpu().numpy() * 255) - np.rint(ref * 255)).max() <= 1
    m = M.LDM(latent_dim=32, pretrained_path=None)
    m.load_state_dict(sd, strict=True)
    m.to(DEV).eval().set_precision(precision)
    eps = torch.randn((b, 32, 16, crop // 8), generator=torch.Generator().manual_seed(9))
    with fixed_randn_like(eps):
        dec, ztd = m.content_style_transfer_wrapper(xc, xs, num_timesteps=steps, eta=0.0)
    assert dec.shape == (b, 1, 128, crop) and ztd.shape == (b, 1, 128, crop)
    # the model part against the oracle on the SAME spectrogram images (so front-end rounding does not enter)
    ref_dec, ref_ztd = O.content_style_transfer(sd, xc.cpu(), xs.cpu(), eps, num_timesteps=steps, eta=0.0, training=False)
    assert_close_mode(dec, ref_dec, precision, "decoded (128x384)")
    assert_close_mode(ztd, ref_ztd, precision, "z_t_decoded (128x384)")
