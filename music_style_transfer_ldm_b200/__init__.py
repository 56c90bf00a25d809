"""mstk: B200-native (sm_100a) denoising hot path of music-style-transfer-ldm.

Public surface mirrors the reference's modules:
    model  (SpectrogramEncoder, SpectrogramDecoder, StyleEncoder, ForwardDiffusion,
            CrossAttention, UNet, SinusoidalPositionEmbeddings, LDM)
    loss   (diffusion_loss, kl_regularization_loss, compression_loss, perceptual_loss,
            VGGishFeatureLoss, style_loss)
    config (config dict)
    train  (LDMTrainer, AutoencoderTrainer, train_autoencoder), audio_processor (AudioPreprocessor),
    dataset (SpectrogramPairDataset, SpectrogramDataset, PackedSpectrograms), autograd (the backward face)
Kernels live in libmstk.so (csrc/, C ABI in include/mstk.h) and are reached through ops.
"""
from . import _lib, ops                      # noqa: F401
from .config import config                   # noqa: F401
from .model import (LDM, CrossAttention, ForwardDiffusion, SinusoidalPositionEmbeddings, SpectrogramDecoder,  # noqa: F401
                    SpectrogramEncoder, StyleEncoder, UNet, set_default_precision)
from .loss import (VGGishFeatureLoss, compression_loss, diffusion_loss, kl_regularization_loss, perceptual_loss,  # noqa: F401
                   style_loss)

__all__ = ["LDM", "CrossAttention", "ForwardDiffusion", "SinusoidalPositionEmbeddings", "SpectrogramDecoder",
           "SpectrogramEncoder", "StyleEncoder", "UNet", "VGGishFeatureLoss", "compression_loss", "diffusion_loss",
           "kl_regularization_loss", "perceptual_loss", "style_loss", "config", "ops", "set_default_precision"]
