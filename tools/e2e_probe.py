"""scratch: why does the pipelined sampling loop slow down at 20 calls?  python tools/e2e_probe.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import music_style_transfer_ldm_b200 as M
from oracle import ldm_oracle as O
import bench
dev = torch.device("cuda", 0)
model = M.LDM(latent_dim=32, pretrained_path=None)
model.load_state_dict(O.make_state_dict(1234), strict=True)
model.to(dev).eval().set_precision("bf16")
g = torch.Generator().manual_seed(100)
ch = torch.rand((64, 1, 128, 128), generator=g).pin_memory()
sh = torch.rand((64, 1, 128, 128), generator=g).pin_memory()
cd, sd_ = ch.to(dev), sh.to(dev)

def loop(k, keep):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    held = []
    for dec, ztd in model.transfer_batches([(ch, sh)] * k, num_timesteps=50, eta=0.0):
        held.append((dec, ztd))
        if len(held) > keep:
            held.pop(0)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k, (time.perf_counter() - t0) * 1e3 / k

with torch.no_grad():
    for _ in range(3):
        model.content_style_transfer_wrapper(cd, sd_, num_timesteps=50, eta=0.0)
    torch.cuda.synchronize()
    print("warm", loop(2, 1))
    for k in (6, 20, 40):
        print("k", k, "no sampler", [round(v, 3) for v in loop(k, 1)])
    s = bench.ClockSampler(0); s.start(); time.sleep(0.3)
    for k in (6, 20, 40):
        print("k", k, "with nvidia-smi -lms 100", [round(v, 3) for v in loop(k, 1)])
    print(s.stop())
    for k in (20,):
        print("k", k, "after sampler", [round(v, 3) for v in loop(k, 1)])
