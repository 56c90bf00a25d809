"""Golden fixture for the autoencoder stage, from the UNMODIFIED reference modules (build container only).

    python tests/golden/make_golden_ae.py      ->  tests/golden/ae_steps.npz

The reference's train_autoencoder (models/train.py:28-138) is one function that also prepares the dataset, needs a
CUDA device and plots; its loop body (train.py:68-83) is driven here verbatim in effect -- the reference's own
SpectrogramEncoder / SpectrogramDecoder / VGGishFeatureLoss / compression_loss objects and the optimizer it
constructs (optim.AdamW over encoder + decoder parameters, train.py:47) -- on CPU, fp32, with seeded inputs.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(os.path.dirname(HERE)), HERE]
import cases      # noqa: E402
import refshim    # noqa: E402
from oracle import ldm_oracle as O   # noqa: E402

STEPS, B, LR = 3, 2, 5e-4
# NOTE on conditioning: d/dz of the KL term's -log(z^2 + 1e-8) (loss.py:31-32) is ~ -1/z, and z is a BatchNorm output, so
# the handful of latent elements nearest zero dominate the encoder's KL gradient and amplify fp32 rounding to ~1e-3 of the
# gradient's max-norm -- in the reference itself.  Adam turns that into sign flips, so from the second update on only the
# LOSSES are comparable tightly; the fixture therefore also stores the step-0 GRADIENTS (the direct check of the backward).
# conv biases feeding a train-mode BatchNorm have a mathematically zero gradient (rounding noise that AdamW's early
# steps amplify to +-lr), so they are not saved; see make_golden_train.py
SAVE_KEYS = ["encoder.encoder.0.weight", "encoder.encoder.1.weight", "encoder.encoder.3.weight", "encoder.encoder.6.weight",
             "encoder.encoder.7.bias", "decoder.decoder.0.weight", "decoder.decoder.4.weight", "decoder.decoder.6.weight",
             "decoder.decoder.6.bias", "encoder.encoder.7.running_var", "decoder.decoder.1.running_mean"]


def step_input(i):
    return torch.rand((B, 1, 128, 128), generator=torch.Generator().manual_seed(9000 + i))


def _sub(sd, prefix):
    return {k[len(prefix):]: v.clone() for k, v in sd.items() if k.startswith(prefix)}


def main():
    torch.manual_seed(0)
    ns = refshim.load()
    sd = O.make_state_dict(cases.WEIGHT_SEED)
    encoder = ns.model.SpectrogramEncoder(32)
    decoder = ns.model.SpectrogramDecoder(32)
    feature_extractor = ns.loss.VGGishFeatureLoss()
    encoder.load_state_dict(_sub(sd, "encoder."), strict=True)
    decoder.load_state_dict(_sub(sd, "decoder."), strict=True)
    feature_extractor.load_state_dict(_sub(sd, "feature_loss_net."), strict=True)
    optimizer = torch.optim.AdamW(list(encoder.parameters()) + list(decoder.parameters()), lr=LR)      # train.py:47
    encoder.train(); decoder.train()
    losses, grads0 = [], {}
    for i in range(STEPS):
        spectrogram = step_input(i)
        latent = encoder(spectrogram)                                                                   # train.py:71-72
        reconstructed = decoder(latent)
        loss = ns.loss.compression_loss(spectrogram, reconstructed, latent, feature_extractor)          # train.py:75
        losses.append(loss.item())
        optimizer.zero_grad()
        loss.backward()
        if i == 0:
            named = {**{"encoder." + k: p for k, p in encoder.named_parameters()}, **{"decoder." + k: p for k, p in decoder.named_parameters()}}
            grads0 = {k: named[k].grad.detach().clone().numpy() for k in SAVE_KEYS if k in named}
        optimizer.step()
        print(i, losses[-1])
    arrays = {"losses": np.asarray(losses, dtype=np.float64)}
    after = {**{"encoder." + k: v for k, v in encoder.state_dict().items()}, **{"decoder." + k: v for k, v in decoder.state_dict().items()}}
    for k in SAVE_KEYS:
        arrays["p:" + k] = after[k].detach().numpy()
    for k, v in grads0.items():
        arrays["g0:" + k] = v
    np.savez_compressed(os.path.join(HERE, "ae_steps.npz"), **arrays)
    print("saved ae_steps.npz")


if __name__ == "__main__":
    main()
