#!/usr/bin/env python
"""Development aid: for every conv/projection of one batch-B UNet step, time the tcgen05 kernel
(graph of 20 back-to-back launches, CUDA events) under the heuristic and under forced (BN, splitk)
choices, and print the clock64 timeline of CTA (0,0,0) for the heuristic choice.

    python tools/tc_tune.py [batch] [--sweep]
stamps: 1 setup done, 2 first A TMA issued, 3 first stage landed, 4 last stage landed, 5 accumulator ready,
        8 partial staged, 9 reduce done+counter, 10 partial reloaded, 11 finalize loop done, 12 store issued,
        6 epilogue done, 7 dealloc
"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_style_transfer_ldm_b200 import ops, _lib

lib = _lib.load()
lib.mstk_debug_tc_timeline.argtypes = [ctypes.c_void_p]
lib.mstk_debug_tc_force.argtypes = [ctypes.c_int] * 4
lib.mstk_debug_tc_last_config.restype = ctypes.POINTER(ctypes.c_int)
B = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 64
SWEEP = "--sweep" in sys.argv
dbg = torch.zeros(16, dtype=torch.int64, device="cuda")

# name, h, w, cin, cout, k, stride, transposed, extras
LAYERS = [("enc1", 16, 16, 32, 64, 3, 1, False, ""), ("enc2", 16, 16, 64, 128, 3, 2, False, "nc"),
          ("enc3", 8, 8, 128, 256, 3, 2, False, ""), ("q2", 4, 4, 256, 256, 1, 1, False, ""),
          ("enc4", 4, 4, 256, 512, 3, 2, False, ""), ("q1", 2, 2, 512, 512, 1, 1, False, ""),
          ("bott", 2, 2, 512, 512, 3, 1, False, ""), ("dec4", 2, 2, 512, 256, 3, 2, True, "add"),
          ("dec3", 4, 4, 256, 128, 3, 2, True, "add"), ("dec2", 8, 8, 128, 64, 3, 2, True, "add"),
          ("dec1", 16, 16, 64, 32, 3, 1, False, "ddim")]


def make(layer):
    name, h, w, cin, cout, k, s, tr, ex = layer
    x = torch.randn((B, h, w, cin), device="cuda").bfloat16()
    wp = (torch.randn((k * k, cout, cin), device="cuda") / (cin * k * k) ** 0.5).bfloat16()
    b = torch.zeros(cout, device="cuda")
    ho, wo = ops.conv_out_hw(h, w, k, s, tr)
    kw = dict(ksize=k, stride=s, transposed=tr, act=1)
    if ex == "nc":
        kw["add_nc"] = torch.randn((B, cout), device="cuda")
    if ex == "add":
        kw["add"] = torch.randn((B, ho, wo, cout), device="cuda").bfloat16()
    if ex == "ddim":
        xc = torch.randn((B, ho, wo, cout), device="cuda")
        kw.update(act=0, out_dtype=torch.float32, want_y=False,
                  ddim=ops.Ddim(x_cur=xc, x_next=xc, x_next_lp=torch.empty_like(xc, dtype=torch.bfloat16), sab_t=0.9, s1m_t=0.4,
                                sab_n=0.91, s1m_n=0.39, eta=0.0))
    flops = 2.0 * B * ho * wo * cout * cin * k * k if not tr else 2.0 * B * h * w * cin * cout * k * k
    return x, wp, b, kw, flops


def time_call(x, wp, b, kw, reps=20):
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(3):
            ops.conv(x, wp, b, **kw)
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for _ in range(reps):
                ops.conv(x, wp, b, **kw)
        g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st); g.replay(); e1.record(st)
        st.synchronize()
    return 1e3 * e0.elapsed_time(e1) / reps


for layer in LAYERS:
    x, wp, b, kw, flops = make(layer)
    lib.mstk_debug_tc_force(0, 0, 0, -1)
    lib.mstk_debug_tc_timeline(None)
    us = time_call(x, wp, b, kw)
    cfg = lib.mstk_debug_tc_last_config()
    cfgs = f"BN={cfg[0]} sk={cfg[1]} st={cfg[2]} smem={cfg[3]//1024}K grid=({cfg[4]},{cfg[5]},{cfg[6]}) KC={cfg[7]}"
    lib.mstk_debug_tc_timeline(dbg.data_ptr())
    dbg.zero_()
    torch.cuda.synchronize()
    ops.conv(x, wp, b, **kw)
    torch.cuda.synchronize()
    t = dbg.cpu().tolist()
    rel = {i: t[i] - t[0] for i in (1, 2, 3, 4, 5, 8, 9, 10, 13, 14, 15, 11, 12, 6, 7) if t[i]}
    lib.mstk_debug_tc_timeline(None)
    print(f"{layer[0]:5s} {us:6.2f} us {flops / us / 1e6:7.1f} TF/s | {cfgs} | " + " ".join(f"{k}:{v}" for k, v in rel.items()), flush=True)
    if SWEEP:
        res = []
        for bn in (16, 32, 64, 128):
            if layer[4] % bn:
                continue
            for sk in (1, 2, 4, 8, 16):
                lib.mstk_debug_tc_force(bn, sk, 0, -1)
                try:
                    u = time_call(x, wp, b, kw)
                except Exception as e:
                    continue
                c = lib.mstk_debug_tc_last_config()
                if c[0] == bn and c[1] == sk:
                    res.append((u, bn, sk, c[2], c[4] * c[5] * c[6]))
        res.sort()
        print("      sweep: " + "  ".join(f"[BN{bn} sk{sk} st{st} ctas{n}: {u:.2f}]" for u, bn, sk, st, n in res[:6]), flush=True)
lib.mstk_debug_tc_force(0, 0, 0, -1)
