"""Golden fixture for the training step, from the UNMODIFIED reference LDMTrainer (build container only).

    python tests/golden/make_golden_train.py      ->  tests/golden/train_steps.npz

Shims beyond refshim.py, none of which touch the arithmetic:
  * torch.autocast -> no-op (fp32 reference; on CPU the reference's autocast would pick bf16);
  * torch.randint / torch.randn_like return the pre-drawn t / noise of the step (train.py:172, model.py:110);
  * the encoder is frozen by hand, as LDM.__init__ does when it loads the pretrained VAE (model.py:340-341;
    pretrained_path=None skips that block because no .pth is obtainable offline).
"""
import contextlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(os.path.dirname(HERE)), HERE]
import cases      # noqa: E402
import refshim    # noqa: E402
from oracle import ldm_oracle as O   # noqa: E402

STEPS, B, LR, W = 3, 2, 5e-4, 3.0
# Not saved on purpose: conv biases that feed a train-mode BatchNorm (decoder.decoder.{0,3}.bias) and the key half of
# in_proj_bias have a mathematically ZERO gradient; what autograd returns for them is rounding noise (~1e-9) whose sign
# Adam turns into full +-lr steps, so their values are not reproducible across summation orders (nor do they matter).
SAVE_KEYS = ["decoder.decoder.6.weight", "decoder.decoder.4.weight", "decoder.decoder.1.weight", "unet.dec1.bias", "unet.enc2.bias",
             "unet.time_mlp.1.bias", "unet.cross_attention2.multihead_attn.in_proj_bias", "style_encoder.enc1.weight",
             "decoder.decoder.1.running_var", "encoder.encoder.1.running_var"]


def step_inputs(i):
    g = torch.Generator().manual_seed(7000 + i)
    return (torch.rand((B, 1, 128, 128), generator=g), torch.rand((B, 1, 128, 128), generator=g),
            torch.randint(0, 200, (B,), generator=g), torch.randn((B, 32, 16, 16), generator=g))


def main():
    torch.manual_seed(0)
    ns = refshim.load(with_train=True)
    sd = O.make_state_dict(cases.WEIGHT_SEED)
    ldm = refshim.build_ldm(sd)
    for p in ldm.encoder.parameters():
        p.requires_grad = False
    tr = ns.train.LDMTrainer(ldm, train_loader=None, device=torch.device("cpu"), lr=LR, style_loss_weight=W)
    ldm.train()
    real_autocast, real_randint, real_randn_like = torch.autocast, torch.randint, torch.randn_like
    losses = []
    try:
        inputs = [step_inputs(i) for i in range(STEPS)]          # drawn before torch.randint / randn_like are patched
        torch.autocast = lambda *a, **k: contextlib.nullcontext()
        for i in range(STEPS):
            x, s, t, eps = inputs[i]
            torch.randint = lambda *a, **k: t.clone()
            torch.randn_like = lambda ref, **k: eps.to(ref.dtype)
            out = tr.train_step(x, s)
            losses.append([out["compression_loss"], out["denoisinsg_loss"], out["style_loss"], out["total_loss"]])
            print(i, out)
    finally:
        torch.autocast, torch.randint, torch.randn_like = real_autocast, real_randint, real_randn_like
    after = ldm.state_dict()
    arrays = {"losses": np.asarray(losses, dtype=np.float64)}
    for k in SAVE_KEYS:
        arrays["p:" + k] = after[k].detach().numpy()
    arrays["sum_all_trainable"] = np.asarray(sum(float(v.double().sum()) for k, v in after.items()
                                                 if k.startswith(("decoder.", "unet.", "style_encoder.")) and v.dtype.is_floating_point))
    np.savez_compressed(os.path.join(HERE, "train_steps.npz"), **arrays)
    print("saved train_steps.npz")


if __name__ == "__main__":
    main()
