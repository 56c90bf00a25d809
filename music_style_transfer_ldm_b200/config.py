"""Hyper-parameters of the hot path: the same flat dict, keys and values as the
reference's models/config.py:2-19, so ``from config import config`` users can switch.

One deliberate difference: ``compression_feature_extractor`` defaults to ``'vggish'``.
The reference default ``'lpips'`` needs the lpips package and AlexNet weights, neither
obtainable offline ("unavailable on this box"); ``'vggish'`` is the
reference's other supported branch (models/loss.py:25-28).
"""

config = dict(
    learning_rate=5e-4,
    learning_rate_factor=0.5,
    learning_rate_patience=5,
    learning_rate_min=1e-6,
    num_epochs=202,
    batch_size=128,
    style_loss_weight=3.0,
    latent_dim_encoder=32,
    data_dir="downloads/",
    processed_spectograms_dataset_folderpath="processed_images",
    pairing_file_path="spectrogram_pair_dataset_pairings.csv",
    unet_num_filters=64,
    forward_diffusion_num_timesteps=200,
    compression_feature_extractor="vggish",
    training_iteration_noise=50,
)
