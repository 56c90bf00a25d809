import sys, os
sys.path[:0] = ["/root/repo", "/root/repo/tests", "/root/repo/tests/golden"]
import torch, numpy as np
import cases
from oracle import ldm_oracle as O, train_oracle as T
import test_gpu_train_autoencoder as TA
import make_golden_ae as GA
sd = O.make_state_dict(cases.WEIGHT_SEED)
ot = T.OracleAutoencoderTrainer(sd, lr=GA.LR)
xs = [torch.rand((4, 1, 128, 128), generator=torch.Generator().manual_seed(9100 + i)) for i in range(20)]
want = np.array([ot.train_step(x) for x in xs])
for prec in ("bf16", "fp32"):
    allr = []
    for rep in range(6):
        enc, dec, tr = TA._trainer(sd, prec, GA.LR)
        got = np.array([tr.train_step(x.cuda()) for x in xs])
        allr.append(np.abs(got - want) / np.abs(want))
    r = np.max(np.array(allr), axis=0)
    print(prec, "max rel diff per step over 6 runs:", " ".join(f"{v:.1e}" for v in r))
