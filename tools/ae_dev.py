"""scratch: per-step loss deviation of AutoencoderTrainer from the oracle trainer, both fp32 routes"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(R, "tests")); sys.path.insert(0, os.path.join(R, "tests", "golden"))
import torch
from oracle import ldm_oracle as O, train_oracle as T
from music_style_transfer_ldm_b200 import ops
import cases
from test_gpu_train_autoencoder import _trainer
sd = O.make_state_dict(cases.WEIGHT_SEED)
for (b, h, w) in [(1, 64, 64), (4, 16, 16), (2, 128, 96)]:
    for tc in (True, False):
        ops.FP32_TC = tc
        enc, dec, tr = _trainer(sd, "fp32", 5e-4)
        ot = T.OracleAutoencoderTrainer(sd, lr=5e-4)
        devs = []
        for i in range(5):
            x = torch.rand((b, 1, h, w), generator=torch.Generator().manual_seed(9300 + 31 * i + b + h + w))
            got, want = tr.train_step(x.cuda()), ot.train_step(x)
            devs.append(abs(got - want) / abs(want))
        print((b, h, w), "FP32_TC" if tc else "cuda-core", " ".join(f"{d:.1e}" for d in devs))
