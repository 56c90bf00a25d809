#!/usr/bin/env python
"""Benchmark of the denoising hot path (BASELINE.json: "UNet denoise steps/s").

    python bench.py --gpus N --steps K --warmup W            # CUDA path (libmstk)
    python bench.py --impl reference --steps K --warmup W    # CPU arm: oracle port on the host cores

Workload (BASELINE config 3): `LDM.content_style_transfer_wrapper(content, style,
num_timesteps=50, eta=0)` at batch 64 per GPU, 128x128 spectrograms (latent 32x16x16):
49 UNet eps-predictions each fused with its DDIM update, plus encoder, style encoder and two
decoder passes.  One bench "step" = one such call.  `value` = UNet denoise steps per second
over the whole job (49 * K * n_gpus / time); one denoise step = one batch-64 UNet evaluation +
DDIM update.  Synthetic inputs, seeded random-init weights (no checkpoint is obtainable offline).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 64
NSTEPS = 50            # "50-step" => 49 UNet evaluations (reference model.py:525)
SPEC = 128
METRIC = "unet_denoise_steps_per_s"
UNIT = "steps/s"


def workload_config(extra=None):
    c = {"workload": "config3: content_style_transfer_wrapper(num_timesteps=50, eta=0), batch 64/GPU, 1x128x128 spectrograms, "
                     "latent 32x16x16, 49 UNet+DDIM steps + encoder + style encoder + 2 decoder passes",
         "batch_per_gpu": BATCH, "num_timesteps": NSTEPS, "unet_evals_per_step": NSTEPS - 1,
         "l2": "flushed between bench steps (256 MiB write); within a step the 13.7 MB of weights are L2 resident by design"}
    if extra:
        c.update(extra)
    return c


# ------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------
# CPU arm (oracle port of the reference's algorithm; the reference tree itself is not on the box)
# ------------------------------------------------------------------------------------------
def cpu_transfer_timing(calls: int, warm_steps: int = 3):
    """Times oracle.content_style_transfer (fp32, eval-mode BN) on the host cores.  Returns
    (denoise steps/s, threads, description)."""
    import torch
    from oracle import ldm_oracle as O
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    threads = torch.get_num_threads()
    sd = O.make_state_dict(1234)
    g = torch.Generator().manual_seed(0)
    x = torch.rand((BATCH, 1, SPEC, SPEC), generator=g)
    s = torch.rand((BATCH, 1, SPEC, SPEC), generator=g)
    eps = torch.randn((BATCH, 32, SPEC // 8, SPEC // 8), generator=g)
    with torch.no_grad():
        O.content_style_transfer(sd, x, s, eps, num_timesteps=warm_steps + 1, eta=0.0, training=False)   # warm-up
        t0 = time.perf_counter()
        for _ in range(calls):
            O.content_style_transfer(sd, x, s, eps, num_timesteps=NSTEPS, eta=0.0, training=False)
        dt = time.perf_counter() - t0
    return (NSTEPS - 1) * calls / dt, threads, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    calls = max(1, args.steps)
    v, threads, dt = cpu_transfer_timing(calls, warm_steps=max(1, args.warmup))
    sample = f"{calls} full config-3 call(s) (batch 64, 49 UNet+DDIM steps, enc + style enc + 2 dec), fp32, torch CPU kernels, {dt:.1f} s"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": calls, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / calls, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config({"arm": "oracle port of the reference algorithm (oracle/ldm_oracle.py) on host cores; "
                                                               "the Python reference tree cannot travel to the GPU box"}),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------
# roofline of the dominant kernel (conv_tc_kernel): per-layer timing of one UNet step
# ------------------------------------------------------------------------------------------
def conv_flops(rec):
    """algorithmic FLOPs of one recorded conv call: 2 * output pixels * cout * cin * taps-per-output"""
    if "flops" in rec:
        return rec["flops"]
    n, h, w, cin, cout, k, s, tr = rec["n"], rec["h"], rec["w"], rec["cin"], rec["cout"], rec["ksize"], rec["stride"], rec["transposed"]
    if tr:
        # every input pixel meets every kernel tap exactly once
        return 2.0 * n * h * w * cin * cout * k * k
    ho, wo = (h + 2 * (k // 2) - k) // s + 1, (w + 2 * (k // 2) - k) // s + 1
    return 2.0 * n * ho * wo * cin * cout * k * k


def kernel_roofline(model, peaks, reps=20):
    """Records the conv calls of ONE UNet+DDIM step, then replays each recorded call `reps` times
    back to back inside a CUDA graph and times the graph with CUDA events on the launch stream.
    Returns the roofline dict for the tcgen05 kernel and a per-layer table."""
    import torch
    from music_style_transfer_ldm_b200 import ops
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(1)
    z = torch.randn((BATCH, 32, 16, 16), generator=g).to(dev)
    st = {"s5": torch.rand((BATCH, 256, 4, 4), generator=g).to(dev), "s6": torch.rand((BATCH, 512, 2, 2), generator=g).to(dev)}
    model.use_cuda_graphs = False
    ops.RECORD = []
    with torch.no_grad():
        model.content_style_ddim_sample(z, st, timesteps=2, eta=0.0)     # exactly one UNet step
    recs, ops.RECORD = ops.RECORD, None
    recs = [r for r in recs if r.get("note") != "hoisted"]      # the style-side projections run once per loop, not per step
    model.use_cuda_graphs = True
    torch.cuda.synchronize()
    rows = []
    stream = torch.cuda.Stream()
    for rec in recs:
        with torch.cuda.stream(stream):
            for _ in range(3):
                rec["replay"]()
            stream.synchronize()
            gr = torch.cuda.CUDAGraph()
            with ops.graph_capture(gr, stream=stream):
                for _ in range(reps):
                    rec["replay"]()
            gr.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); gr.replay(); e1.record(stream)
            stream.synchronize()
            us = 1e3 * e0.elapsed_time(e1) / reps
        rows.append({"layer": rec["tag"], "kernel": rec["kernel"], "us": us, "gflop": conv_flops(rec) / 1e9,
                     "tflops": conv_flops(rec) / (us * 1e-6) / 1e12})
    tc = [r for r in rows if r["kernel"] == "conv_tc_kernel"]
    fl, us = sum(r["gflop"] for r in tc) * 1e9, sum(r["us"] for r in tc)
    peak = peaks.get("bf16_tflops", 1590.0)
    achieved = fl / (us * 1e-6) / 1e12 if us > 0 else 0.0
    # the same step as ONE graph (what the sampling loop replays): PDL overlap = 1 - step / sum of isolated launches
    with torch.cuda.stream(stream):
        for r in recs:
            r["replay"]()
        stream.synchronize()
        gr = torch.cuda.CUDAGraph()
        with ops.graph_capture(gr, stream=stream):
            for _ in range(reps):
                for r in recs:
                    r["replay"]()
        gr.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); gr.replay(); e1.record(stream)
        stream.synchronize()
        chain_us = 1e3 * e0.elapsed_time(e1) / reps
    iso_us = sum(r["us"] for r in rows)
    traffic, traffic_src = None, None
    try:        # dram bytes per launch from the committed ncu --set full capture (tools/traffic_from_ncu.py)
        with open(os.path.join(ROOT, "profiles", "conv_tc_traffic.json")) as f:
            tj = json.load(f)
        traffic, traffic_src = tj["bytes_per_launch"], f"profiles/conv_tc_traffic.json ({tj['launches']} launches of one UNet step, {tj['metric']})"
    except (OSError, KeyError, ValueError):
        pass
    roof = {"bound": "tensor", "kernel": "conv_tc_kernel", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
            "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "launches_per_unet_step": len(tc),
            "flop_per_launch_avg": fl / max(1, len(tc)), "us_per_launch_avg": us / max(1, len(tc)),
            "launches_per_unet_step_all_kernels": len(rows), "sum_isolated_us_all_kernels": iso_us, "chained_step_us": chain_us,
            "pdl_overlap_frac": 1.0 - chain_us / iso_us if iso_us > 0 else None,
            "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst)" if "bf16_tflops" in peaks else "fallback 1590 (B200_PROFILING.md)",
            "note": "achieved = sum of algorithmic conv FLOPs of the tcgen05 launches of one batch-64 UNet step / sum of their "
                    "isolated launch durations (each layer replayed 20x in a CUDA graph, CUDA events)"}
    return roof, rows


# ------------------------------------------------------------------------------------------
# VAE leg (BASELINE config 2): encode -> decode round trip of 16 spectrograms of 128 mel x 1024 frames; HBM bound
# ------------------------------------------------------------------------------------------
VAE_BATCH, VAE_H, VAE_W = 16, 128, 1024
VAE_ELEMS_PER_SAMPLE = 12.98e6        # input + every inter-layer activation written once and read once + output


def vae_leg(sd, dev, peaks, precision, reps=20):
    """SpectrogramEncoder -> SpectrogramDecoder through the public modules (NCHW fp32 in / out), captured once in a
    CUDA graph and replayed; eval mode (BatchNorm folded into the convs) and train mode (batch statistics in the conv
    epilogue + one normalise pass).  achieved = SURVEY 8(d)'s minimal traffic / device time."""
    import torch
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200 import ops
    enc, dec = M.SpectrogramEncoder(32), M.SpectrogramDecoder(32)
    enc.load_state_dict({k[len("encoder."):]: v for k, v in sd.items() if k.startswith("encoder.")})
    dec.load_state_dict({k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")})
    enc.to(dev).set_precision(precision); dec.to(dev).set_precision(precision)
    x = torch.rand((VAE_BATCH, 1, VAE_H, VAE_W), generator=torch.Generator().manual_seed(3)).to(dev)
    bytes_min = VAE_ELEMS_PER_SAMPLE * VAE_BATCH * (2 if precision == "bf16" else 4)
    peak = peaks.get("hbm_gbs", 6550.0)
    out = {"workload": f"config2: decoder(encoder(x)), x [16,1,128,1024] fp32 NCHW, {precision} activations, latent [16,32,16,128]",
           "bound": "hbm", "peak": peak, "unit": "GB/s", "algorithmic_bytes_per_round_trip": bytes_min,
           "l2": "working set of one round trip (>= 415 MB bf16) exceeds the 126 MB L2",
           "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6550 (B200_PROFILING.md)"}
    with torch.no_grad():
        for mode in ("eval", "train"):
            enc.train(mode == "train"); dec.train(mode == "train")
            stream = torch.cuda.Stream()
            with torch.cuda.stream(stream):
                for _ in range(3):
                    y = dec(enc(x))
                stream.synchronize()
                gr = torch.cuda.CUDAGraph()
                with ops.graph_capture(gr, stream=stream):
                    y = dec(enc(x))
                gr.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(reps):
                    gr.replay()
                e1.record(stream)
                stream.synchronize()
            us = 1e3 * e0.elapsed_time(e1) / reps
            out[mode] = {"us_per_round_trip": us, "spectrograms_per_s": VAE_BATCH / (us * 1e-6), "achieved": bytes_min / (us * 1e-6) / 1e9,
                         "frac": bytes_min / (us * 1e-6) / 1e9 / peak, "finite": bool(torch.isfinite(y).all())}
    return out


# ------------------------------------------------------------------------------------------
# training leg (BASELINE config 4): LDMTrainer.train_step, bf16, GLOBAL batch 256 sharded over the ranks
# ------------------------------------------------------------------------------------------
TRAIN_GLOBAL_BATCH = 256


def train_leg(args, rank, world, dev, barrier, global_batch=None, rooflines=True):
    """Returns a dict for the JSON line (rank 0) with train samples/s: device-resident and end-to-end
    (pinned host batch -> device inside the timed region, 4 loss scalars read back every step)."""
    import torch
    import torch.distributed as dist
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200 import ops
    from music_style_transfer_ldm_b200.train import LDMTrainer
    from oracle import ldm_oracle as O
    gb = global_batch or TRAIN_GLOBAL_BATCH
    if gb % world:
        return {"skipped": f"global batch {gb} not divisible by {world} ranks"}
    per = gb // world
    model = M.LDM(latent_dim=32, pretrained_path=None)
    model.load_state_dict(O.make_state_dict(1234), strict=True)
    model._freeze_encoder()
    model.to(dev).set_precision(args.precision)
    tr = LDMTrainer(model, None, dev, lr=5e-4, style_loss_weight=3.0, process_group=None, world_size=world)
    model.train()
    g = torch.Generator().manual_seed(500 + rank)
    xh = torch.rand((per, 1, SPEC, SPEC), generator=g).pin_memory()
    sh = torch.rand((per, 1, SPEC, SPEC), generator=g).pin_memory()
    xd, sd_ = xh.to(dev), sh.to(dev)
    loss_h = torch.empty(4).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    steps = max(2, min(args.steps, 8))

    def step_device():
        return tr.train_step_device(xd, sd_)

    def step_e2e():
        out = tr.train_step_device(xh.to(dev, non_blocking=True), sh.to(dev, non_blocking=True))
        loss_h.copy_(torch.stack([out[k] for k in ("compression_loss", "denoisinsg_loss", "style_loss", "total_loss")]), non_blocking=True)

    def timed(fn, k):
        evs = []
        for _ in range(k):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    def loop_e2e(k):
        """k steps the way LDMTrainer.train_epoch feeds them (train.device_prefetch: the pinned-host -> device copy of batch
        i+1 is enqueued on a copy stream under the kernels of step i), the four losses read back every step; one event pair
        around the whole loop, the pipeline's unoverlapped first upload included"""
        import gc
        from music_style_transfer_ldm_b200.train import device_prefetch
        gc.collect()
        gc.disable()                                         # no generation-2 pause of the host inside the loop
        try:
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for c, s in device_prefetch([((xh, None), (sh, None))] * k, dev, pick=lambda el: (el[0][0], el[1][0])):
                out = tr.train_step_device(c, s)
                loss_h.copy_(torch.stack([out[q] for q in ("compression_loss", "denoisinsg_loss", "style_loss", "total_loss")]), non_blocking=True)
            e1.record()
            torch.cuda.synchronize()
        finally:
            gc.enable()
        return e0.elapsed_time(e1)

    for _ in range(3):
        step_device()
    step_e2e()
    loop_e2e(2)
    torch.cuda.synchronize()
    n0 = ops.LAUNCHES
    step_device()
    torch.cuda.synchronize()
    launches = ops.LAUNCHES - n0
    barrier()
    ms = timed(step_device, steps)
    barrier()
    ms_serial = timed(step_e2e, steps)
    barrier()
    e2e_runs = [loop_e2e(steps), loop_e2e(steps)]            # host-paced loop: the faster of two runs, both printed
    ms_e2e = min(e2e_runs)
    barrier()
    t = torch.tensor([ms, ms_e2e, ms_serial], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_serial = t.tolist()
    final = [float(v) for v in loss_h.tolist()]
    extra = {}
    if rank == 0 and rooflines:
        extra = train_rooflines(model, per, dev)
    tr.close()
    del tr, model
    torch.cuda.empty_cache()
    return {"metric": "train_samples_per_s", "value": gb * steps / (ms * 1e-3), "unit": "samples/s",
            "ms_per_step": ms / steps, "steps": steps, "global_batch": gb, "batch_per_gpu": per,
            "scaling": "strong" if gb == TRAIN_GLOBAL_BATCH else "weak",
            "e2e": {"value": gb * steps / (ms_e2e * 1e-3), "unit": "samples/s", "ms_per_step": ms_e2e / steps,
                    "h2d_bytes_per_step": 2 * xh.numel() * 4, "d2h_bytes_per_step": 16,
                    "how": "train_epoch's feed: upload of batch i+1 on a copy stream under step i; one event pair around all steps, "
                           "no L2 flush between steps (a step streams > 10 GB, the L2 holds 126 MB)",
                    "runs_ms_per_step": [r / steps for r in e2e_runs],
                    "serial_value": gb * steps / (ms_serial * 1e-3), "serial_ms_per_step": ms_serial / steps},
            "gpu_launches_per_step": int(launches), "last_losses": final, **extra,
            "workload": "config4: LDMTrainer.train_step (fwd + 4 losses incl. 2 VGGish feature losses + bwd + Adam), bf16, "
                        "1x128x128 spectrograms, encoder frozen, style_loss_weight 3.0, DDP gradient + BatchNorm-stat all-reduce"}


def ae_leg(sd, dev, precision, batch=256, steps=6):
    """one optimisation step of the autoencoder stage (train.py:68-83): encoder + decoder forward with
    train-mode BatchNorm, compression_loss (MSE + 0.1 VGGish + 0.01 KL), backward through both, fused AdamW; CUDA-graph replay."""
    import torch
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200 import ops
    from music_style_transfer_ldm_b200.loss import VGGishFeatureLoss
    from music_style_transfer_ldm_b200.train import AutoencoderTrainer
    enc, dec, fl = M.SpectrogramEncoder(32), M.SpectrogramDecoder(32), VGGishFeatureLoss()
    enc.load_state_dict({k[len("encoder."):]: v for k, v in sd.items() if k.startswith("encoder.")})
    dec.load_state_dict({k[len("decoder."):]: v for k, v in sd.items() if k.startswith("decoder.")})
    fl.load_state_dict({k[len("feature_loss_net."):]: v for k, v in sd.items() if k.startswith("feature_loss_net.")})
    enc.to(dev).set_precision(precision); dec.to(dev).set_precision(precision); fl.to(dev); fl.precision = precision
    tr = AutoencoderTrainer(enc, dec, fl, dev, lr=5e-4)
    enc.train(); dec.train()
    xh = torch.rand((batch, 1, SPEC, SPEC), generator=torch.Generator().manual_seed(77)).pin_memory()
    xd = xh.to(dev)
    loss_h = torch.empty(1).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step_device():
        return tr.train_step_device(xd)

    def step_e2e():
        loss_h.copy_(tr.train_step_device(xh.to(dev, non_blocking=True)).reshape(1), non_blocking=True)

    def timed(fn, k):
        evs = []
        for _ in range(k):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    for _ in range(4):
        step_device()
    step_e2e()
    torch.cuda.synchronize()
    n0 = ops.LAUNCHES
    step_device()
    torch.cuda.synchronize()
    launches = ops.LAUNCHES - n0
    ms, ms_e2e = timed(step_device, steps), timed(step_e2e, steps)
    return {"metric": "autoencoder_train_samples_per_s", "value": batch * steps / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms / steps,
            "steps": steps, "batch": batch, "e2e": {"value": batch * steps / (ms_e2e * 1e-3), "unit": "samples/s", "ms_per_step": ms_e2e / steps,
                                                    "h2d_bytes_per_step": xh.numel() * 4, "d2h_bytes_per_step": 4},
            "gpu_launches_per_step": int(launches), "last_loss": float(loss_h[0]),
            "workload": f"train_autoencoder step (models/train.py:68-83), {precision}, batch {batch}, 1x128x128 spectrograms, 'vggish' perceptual branch"}


def _time_us(fn, reps=10):
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / reps


def train_rooflines(model, per, dev):
    """Rooflines of the training step's two kernel families, measured live (CUDA events, each call repeated
    back to back, inputs far larger than L2 for the big layers):
      * tensor: the tcgen05 convolutions of the VGGish feature stack (17 of the step's 19.6 GFLOP/sample), per layer
        and in aggregate, against the measured bf16 peak;
      * hbm: the streaming kernels of the feature loss on the first (largest) tap, against the measured copy bandwidth.
    Algorithmic work: conv = 2*pixels*cout*cin*9 FLOP; sample_stats = one read of the map; feature_mse3 = one read of
    the three maps; maxpool = one read + a quarter-size write."""
    import torch
    from music_style_transfer_ldm_b200 import ops
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    g = torch.Generator().manual_seed(3)
    three = torch.rand((3 * per, SPEC, SPEC, 1), generator=g).to(dev)
    fl = model.feature_loss_net
    ops.RECORD = []
    with torch.no_grad():
        taps = fl._taps(three)
    recs, ops.RECORD = ops.RECORD, None
    rows = []
    for rec in recs:
        if rec["kernel"] != "conv_tc_kernel":
            continue
        us = _time_us(rec["replay"], 5)
        rows.append({"layer": rec["tag"], "us": us, "gflop": conv_flops(rec) / 1e9, "tflops": conv_flops(rec) / (us * 1e-6) / 1e12})
    fl_sum, us_sum = sum(r["gflop"] for r in rows) * 1e9, sum(r["us"] for r in rows)
    peak_t = peaks.get("bf16_tflops", 1590.0)
    tensor = {"bound": "tensor", "kernel": "conv_tc_kernel (VGGish stack, 3 x %d images)" % per, "achieved": fl_sum / (us_sum * 1e-6) / 1e12,
              "peak": peak_t, "unit": "TFLOP/s", "frac": fl_sum / (us_sum * 1e-6) / 1e12 / peak_t, "layers": rows,
              "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst)" if "bf16_tflops" in peaks else "fallback 1590"}
    f0 = taps[0]
    nbytes = f0.numel() * f0.element_size()
    peak_h = peaks.get("hbm_gbs", 6500.0)
    stats = torch.zeros(6 * per + 4, dtype=torch.float64, device=dev)
    oa, oc = torch.zeros(1, device=dev), torch.zeros(1, device=dev)
    lib = ops._lib.load()

    def k_stats():
        lib.mstk_sample_stats(f0.data_ptr(), ops.dtype_code(f0), 3 * per, f0.numel() // (3 * per), stats.data_ptr(), ops._stream())

    us_stats = _time_us(k_stats)
    stats.zero_()
    us_mse = _time_us(lambda: ops.feature_mse3(f0, per, 1.0, oa, oc)) - us_stats          # feature_mse3 = stats + mse kernels
    us_pool = _time_us(lambda: ops.maxpool2x2(f0))
    hbm = [{"kernel": "sample_stats_kernel", "bytes": nbytes, "us": us_stats},
           {"kernel": "feat_mse3_kernel", "bytes": nbytes, "us": us_mse},
           {"kernel": "maxpool2x2_vec_kernel", "bytes": nbytes * 1.25, "us": us_pool}]
    for h in hbm:
        h["achieved_gbs"] = h["bytes"] / (h["us"] * 1e-6) / 1e9
        h["frac"] = h["achieved_gbs"] / peak_h
    return {"roofline": tensor, "roofline_hbm": {"bound": "hbm", "peak": peak_h, "unit": "GB/s", "tap": "VGGish tap 1: [%d,128,128,64] bf16" % (3 * per),
                                                 "kernels": hbm}}


# ------------------------------------------------------------------------------------------
# the bar on the same box: stock PyTorch (cuDNN / cuBLAS) running the reference's algorithm on the B200
# ------------------------------------------------------------------------------------------
def gpu_baseline(sd, dev, quick=False):
    """SURVEY 8(d) / BASELINE.md section 3: the reference ships no kernels of its own, so the existing-GPU-implementation
    bar is stock PyTorch executing the same modules on this B200.  The functional restatement (oracle/ldm_oracle.py,
    pinned to the reference) is run with its tensors on `cuda`: (i) eager fp32 with torch's TF32 defaults, (ii) under
    torch.autocast(bfloat16), (iii) (ii) captured in one torch.cuda.CUDAGraph.  Same config-3 / config-4 / config-2
    inputs as the product legs, CUDA events, rank 0 only.  This measures the bar; nothing here is product code."""
    import torch
    from oracle import ldm_oracle as O, train_oracle as T
    sdc = {k: v.to(dev) for k, v in sd.items()}
    g = torch.Generator().manual_seed(0)
    x = torch.rand((BATCH, 1, SPEC, SPEC), generator=g).to(dev)
    s = torch.rand((BATCH, 1, SPEC, SPEC), generator=g).to(dev)
    eps = torch.randn((BATCH, 32, SPEC // 8, SPEC // 8), generator=g).to(dev)
    out = {"what": "oracle/ldm_oracle.py (functional torch restatement of models/model.py) on cuda: cuDNN / cuBLAS kernels of torch "
                   + torch.__version__, "unit_sampling": UNIT, "unit_train": "samples/s"}

    def time_ms(fn, reps):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    def call3():
        return O.content_style_transfer(sdc, x, s, eps, num_timesteps=NSTEPS, eta=0.0, training=False)

    def call3_bf16():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return call3()

    reps = 2 if quick else 5
    with torch.no_grad():
        ms = time_ms(call3, reps)
        out["config3_eager_fp32_tf32"] = {"value": (NSTEPS - 1) / (ms * 1e-3), "ms_per_call": ms}
        ms = time_ms(call3_bf16, reps)
        out["config3_autocast_bf16"] = {"value": (NSTEPS - 1) / (ms * 1e-3), "ms_per_call": ms}
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                call3_bf16()
                st.synchronize()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=st):
                    call3_bf16()
                st.synchronize()
            ms = time_ms(gr.replay, reps)
            out["config3_autocast_bf16_cudagraph"] = {"value": (NSTEPS - 1) / (ms * 1e-3), "ms_per_call": ms}
            del gr
        except Exception as e:      # noqa: BLE001 -- a baseline that cannot be captured is reported, not fatal
            out["config3_autocast_bf16_cudagraph"] = {"error": repr(e)[:200]}
        # config 2: VAE round trip, eval-mode BatchNorm
        xv = torch.rand((VAE_BATCH, 1, VAE_H, VAE_W), generator=torch.Generator().manual_seed(3)).to(dev)

        def vae():
            return O.decoder_fwd(sdc, O.encoder_fwd(sdc, xv, False), False)

        def vae_bf16():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return vae()

        out["config2_eval_eager_fp32_tf32"] = {"us_per_round_trip": 1e3 * time_ms(vae, reps)}
        out["config2_eval_autocast_bf16"] = {"us_per_round_trip": 1e3 * time_ms(vae_bf16, reps)}
    # config 4: one training step at global batch 256 (fwd + 4 losses + autograd backward + torch.optim.Adam)
    gb = TRAIN_GLOBAL_BATCH
    xt = torch.rand((gb, 1, SPEC, SPEC), generator=g).to(dev)
    stl = torch.rand((gb, 1, SPEC, SPEC), generator=g).to(dev)
    tt = torch.randint(0, 200, (gb,), generator=g).to(dev)
    ee = torch.randn((gb, 32, SPEC // 8, SPEC // 8), generator=g).to(dev)
    for name, ac in (("config4_eager_fp32_tf32", False), ("config4_autocast_bf16", True)):
        try:
            tr = T.OracleTrainer(sdc, lr=5e-4, style_loss_weight=3.0)

            def step():
                if ac:
                    with torch.autocast("cuda", dtype=torch.bfloat16):
                        tr.train_step(xt, stl, tt, ee)
                else:
                    tr.train_step(xt, stl, tt, ee)

            ms = time_ms(step, 2 if quick else 3)
            out[name] = {"value": gb / (ms * 1e-3), "ms_per_step": ms}
            del tr
        except Exception as e:      # noqa: BLE001
            out[name] = {"error": repr(e)[:200]}
        torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------
# config 5: waveforms -> mel front end -> 50-step transfer (per GPU: 64 content + 64 style clips of 10 s)
# ------------------------------------------------------------------------------------------
MEL_CLIPS, MEL_LEN, MEL_CROP = 64, 220500, 384
MEL_BYTES_PER_CLIP = MEL_LEN * 4 + 128 * 431 * 4        # 882 KB in + 221 KB fp32 out = 1.10 MB


def mel_leg(model, dev, peaks, rank, cpu_baseline=True, reps=10):
    """BASELINE config 5 on this rank: 0.1*randn(64, 220500) content and style clips -> AudioPreprocessor.spectrogram_batch
    (csrc/mel.cu, crop 128x384) -> content_style_transfer_wrapper(num_timesteps=50).  Reports the mel front end alone (HBM
    roofline: 1.10 MB per clip) and the whole config-5 call (UNet denoise steps/s at latent 16x48), device time."""
    import torch
    from music_style_transfer_ldm_b200.audio_processor import AudioPreprocessor
    g = torch.Generator().manual_seed(900 + rank)
    wc = (0.1 * torch.randn((MEL_CLIPS, MEL_LEN), generator=g)).to(dev)
    ws = (0.1 * torch.randn((MEL_CLIPS, MEL_LEN), generator=g)).to(dev)
    ap = AudioPreprocessor()

    def mel_only():
        return ap.spectrogram_batch(wc, n_mels=128, crop_frames=MEL_CROP)

    def full():
        xc = ap.spectrogram_batch(wc, n_mels=128, crop_frames=MEL_CROP)
        xs = ap.spectrogram_batch(ws, n_mels=128, crop_frames=MEL_CROP)
        return model.content_style_transfer_wrapper(xc, xs, num_timesteps=NSTEPS, eta=0.0)

    with torch.no_grad():
        us_mel = _time_us(mel_only, reps)
        us_full = _time_us(full, max(2, reps // 3))
    peak = peaks.get("hbm_gbs", 6550.0)
    gbs = MEL_CLIPS * MEL_BYTES_PER_CLIP / (us_mel * 1e-6) / 1e9
    out = {"workload": "config5 (per GPU): 2 x 64 clips of 10 s @ 22.05 kHz -> mel 128x431 -> crop 128x384 -> [0,1] -> 50-step transfer "
                       "(latent 32x16x48) -> decode",
           "mel": {"kernel": "mel front end (csrc/mel.cu)", "clips": MEL_CLIPS, "us": us_mel, "clips_per_s": MEL_CLIPS / (us_mel * 1e-6),
                   "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                                "algorithmic_bytes_per_clip": MEL_BYTES_PER_CLIP}},
           "end_to_end": {"ms_per_call": us_full * 1e-3, "unet_denoise_steps_per_s": (NSTEPS - 1) / (us_full * 1e-6),
                          "clips_per_s": MEL_CLIPS / (us_full * 1e-6)}}
    # the step after the path: 64 decoded 128 x 128 images -> NNLS magnitude -> 32 Griffin-Lim iterations -> waveforms
    from music_style_transfer_ldm_b200 import ops as _ops
    img = (torch.rand((MEL_CLIPS, 128, 128), generator=g) * 255).to(torch.uint8).to(dev)
    u = torch.rand((MEL_CLIPS, 128, 1025), generator=g).to(dev)
    mag = ap.mel_images_to_magnitude(img)
    with torch.no_grad():
        us_inv = _time_us(lambda: ap.audio_batch(img, phase_u=u), max(2, reps // 3))
        us_gl0 = _time_us(lambda: _ops.griffinlim(mag, u, n_iter=0), max(2, reps // 3))
        us_gl32 = _time_us(lambda: _ops.griffinlim(mag, u, n_iter=32), max(2, reps // 3))
    us_it = (us_gl32 - us_gl0) / 32.0
    # per frame and iteration: angles read + written, tprev read + written (complex64), S read, windowed frame written + read (2048 fp32)
    gl_bytes_it = MEL_CLIPS * 128 * (1025 * (4 * 8 + 4) + 2 * 2048 * 4)
    out["inverse"] = {"workload": "64 uint8 mel images 128 x 128 -> mel_to_stft (clipped pseudo-inverse = librosa's NNLS result on a dB image) -> "
                                  "32 Griffin-Lim iterations (n_fft 2048, hop 512) -> 64 x 65 024 samples (audio_processor.py:81-100)",
                      "kernel": "gl_istft_kernel + gl_stft_kernel (csrc/griffinlim.cu)", "us_per_call": us_inv, "clips_per_s": MEL_CLIPS / (us_inv * 1e-6),
                      "us_per_iteration": us_it,
                      "roofline": {"bound": "hbm", "achieved": gl_bytes_it / (us_it * 1e-6) / 1e9, "peak": peak, "unit": "GB/s",
                                   "frac": gl_bytes_it / (us_it * 1e-6) / 1e9 / peak, "algorithmic_bytes_per_iteration": gl_bytes_it}}
    if cpu_baseline:
        from oracle import audio_inverse_oracle as AO
        t0 = time.perf_counter()
        AO.grayscale_mel_spectogram_image_to_audio(img[0].cpu().numpy(), u[0].cpu().numpy().T.astype("float64"), use_lbfgs=False)
        dt = time.perf_counter() - t0
        out["inverse"]["cpu_baseline"] = {"clips_per_s": 1.0 / dt, "kind": "port", "cores": 1,
                                          "sample": f"1 clip through oracle/audio_inverse_oracle.py (numpy restatement of librosa's mel_to_audio), {dt:.2f} s"}
    if cpu_baseline:
        from oracle import mel_oracle as MO
        w = wc[:4].cpu().numpy()
        t0 = time.perf_counter()
        for i in range(w.shape[0]):
            MO.waveform_to_unit_image(w[i], 22050, 128, MEL_CROP)
        dt = time.perf_counter() - t0
        out["mel"]["cpu_baseline"] = {"clips_per_s": w.shape[0] / dt, "kind": "port", "cores": 1,
                                      "sample": f"{w.shape[0]} clips through oracle/mel_oracle.py (numpy restatement of librosa's defaults; librosa "
                                                f"is not installable here), {dt:.2f} s"}
    return out


def fp32_leg(sd, dev, content_d, style_d, calls=3):
    """the same config-3 call in fp32 mode (1e-4 parity mode): UNet denoise steps/s, device time"""
    import torch
    import music_style_transfer_ldm_b200 as M
    m = M.LDM(latent_dim=32, pretrained_path=None)
    m.load_state_dict(sd, strict=True)
    m.to(dev).eval().set_precision("fp32")
    with torch.no_grad():
        us = _time_us(lambda: m.content_style_transfer_wrapper(content_d, style_d, num_timesteps=NSTEPS, eta=0.0), calls)
    del m
    torch.cuda.empty_cache()
    return {"metric": METRIC, "value": (NSTEPS - 1) / (us * 1e-6), "unit": UNIT, "ms_per_call": us * 1e-3, "precision": "fp32",
            "unet_step_us": us / (NSTEPS - 1)}


def batch_sweep(model, dev, peaks, flop_per_sample_step, batches=(64, 128, 256, 512), calls=3):
    """Not a BASELINE config: the config-3 call at larger batches on the same kernels and graph -- how much of the 0.06 is the
    batch size of the config (11 dependent launches of 0.1-0.9 waves per step at batch 64) and how much the kernels.  Device
    time of the whole call per UNet step, sample-steps/s, fraction of the measured bf16 peak on the algorithmic conv FLOPs."""
    import torch
    peak = float(peaks.get("bf16_tflops", 1590.0))
    out = []
    for b in batches:
        g = torch.Generator().manual_seed(900 + b)
        c, s_ = torch.rand((b, 1, SPEC, SPEC), generator=g).to(dev), torch.rand((b, 1, SPEC, SPEC), generator=g).to(dev)
        with torch.no_grad():
            us = _time_us(lambda: model.content_style_transfer_wrapper(c, s_, num_timesteps=NSTEPS, eta=0.0), calls)
        step_us = us / (NSTEPS - 1)
        tf = flop_per_sample_step * b / (step_us * 1e-6) / 1e12
        out.append({"batch": b, "unet_step_us": step_us, "sample_steps_per_s": b * (NSTEPS - 1) / (us * 1e-6), "tflops": tf, "frac_of_bf16_peak": tf / peak})
        del c, s_
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------
# data-parallel self-check (the driver's test box has one GPU: the N > 1 bench runs carry the evidence)
# ------------------------------------------------------------------------------------------
def ddp_parity(args, rank, world, dev):
    """3 steps of the SHARDED training step at global batch 256 (graph warm-up 1: eager, capture, replay -- so the captured
    DDP step with its in-graph NCCL all-reduces is covered) against (a) the same global batch run single-process on rank 0
    and (b) the fp32 oracle curve tests/golden/train_curve_b256.npz.  Also the spread of the replicas' parameter checksums."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200.train import LDMTrainer
    from oracle import ldm_oracle as O
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden_curve256 as G256
    keys = ("compression_loss", "denoisinsg_loss", "style_loss", "total_loss")
    per = G256.B // world

    def run(w, shard):
        model = M.LDM(latent_dim=32, pretrained_path=None)
        model.load_state_dict(O.make_state_dict(1234), strict=True)
        model._freeze_encoder()
        model.to(dev).set_precision(args.precision)
        tr = LDMTrainer(model, None, dev, lr=G256.LR, style_loss_weight=G256.W, process_group=None, world_size=w)
        tr.graph_warmup = 1
        model.train()
        rows = []
        for i in range(G256.STEPS):
            x, s, t, eps = (v[shard].to(dev) for v in G256.step_inputs(i))
            out = tr.train_step_device(x, s, t=t, noise=eps)
            rows.append([float(out[k]) for k in keys])
        chk = tr.arena.p.double().sum().reshape(1)
        tr.close()
        del tr, model
        torch.cuda.empty_cache()
        return np.asarray(rows), chk

    ddp, chk = run(world, slice(rank * per, (rank + 1) * per))
    chks = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(chks, chk)
    res = None
    if rank == 0:
        single, _ = run(1, slice(0, G256.B))
        gold = np.load(os.path.join(ROOT, "tests", "golden", "train_curve_b256.npz"))["losses"]
        c = torch.cat(chks).cpu().numpy()
        res = {"steps": int(G256.STEPS), "global_batch": int(G256.B), "path": "step 0 eager, step 1 graph capture, step 2 graph replay",
               "max_rel_loss_dev": float(np.max(np.abs(ddp - single) / np.abs(single))),
               "max_rel_loss_dev_vs_oracle_fixture": float(np.max(np.abs(ddp - gold) / np.abs(gold))),
               "replica_checksum_spread": float((c.max() - c.min()) / max(1e-30, abs(c[0]))), "tolerance": 1e-2}
        res["pass"] = bool(res["max_rel_loss_dev"] <= 1e-2 and res["max_rel_loss_dev_vs_oracle_fixture"] <= 1e-2
                           and res["replica_checksum_spread"] <= 1e-6)
    dist.barrier()
    return res


# ------------------------------------------------------------------------------------------
# CUDA arm
# ------------------------------------------------------------------------------------------
def run_cuda(args):
    import torch
    import torch.distributed as dist
    import music_style_transfer_ldm_b200 as M
    from music_style_transfer_ldm_b200 import ops
    from oracle import ldm_oracle as O      # only for the shared seeded weights and the cpu_baseline leg

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sd = O.make_state_dict(1234)
    model = M.LDM(latent_dim=32, pretrained_path=None)
    model.load_state_dict(sd, strict=True)
    model.to(dev).eval().set_precision(args.precision)

    g = torch.Generator().manual_seed(100 + rank)            # every rank: its own shard of the global batch
    content_h = torch.rand((BATCH, 1, SPEC, SPEC), generator=g).pin_memory()
    style_h = torch.rand((BATCH, 1, SPEC, SPEC), generator=g).pin_memory()
    content_d, style_d = content_h.to(dev), style_h.to(dev)
    out_h = [torch.empty((BATCH, 1, SPEC, SPEC)).pin_memory() for _ in range(2)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step_device():
        return model.content_style_transfer_wrapper(content_d, style_d, num_timesteps=NSTEPS, eta=0.0)

    def step_e2e():
        c = content_h.to(dev, non_blocking=True)
        s = style_h.to(dev, non_blocking=True)
        dec, ztd = model.content_style_transfer_wrapper(c, s, num_timesteps=NSTEPS, eta=0.0)
        out_h[0].copy_(dec, non_blocking=True)
        out_h[1].copy_(ztd, non_blocking=True)

    def timed(fn, k):
        evs = []
        for _ in range(k):
            flush.fill_(1)                                   # L2 flush, outside the timed events
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)        # ms on the device

    def loop_e2e(k):
        """k calls through LDM.transfer_batches: pinned host batches in, pinned host images out, the upload of call i+1 and
        the download of call i-1 on copy streams under call i.  One event pair around the whole loop (first upload and last
        download unoverlapped, included); the host waits for the last download before the closing event is recorded."""
        import gc
        gc.collect()
        gc.disable()                                         # a generation-2 collection is a 30-80 ms host pause: longer than the queued work
        try:
            def feed():
                for _ in range(k):
                    # L2 flush (144 MiB > the 126 MB L2) INSIDE the timed region: the loop pulls batch i+1 just before it
                    # enqueues call i, so exactly one flush sits between consecutive calls on the compute stream
                    flush[:144 << 20].fill_(1)
                    yield content_h, style_h

            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for dec, ztd in model.transfer_batches(feed(), num_timesteps=NSTEPS, eta=0.0, ring=8, depth=6):
                out_h[0], out_h[1] = dec, ztd
            e1.record()
            torch.cuda.synchronize()
        finally:
            gc.enable()
        return e0.elapsed_time(e1)

    with torch.no_grad():
        for _ in range(max(3, args.warmup)):
            step_device()
        step_e2e()
        loop_e2e(10)                                         # fills the pinned-buffer cache (cudaHostAlloc of the ring's 16 tensors is ~25 ms) outside the timing
        torch.cuda.synchronize()
        ops.LAUNCHES = 0
        step_device()
        torch.cuda.synchronize()
        launches_per_step = ops.LAUNCHES

        sampler = ClockSampler(local)
        if rank == 0 and not os.environ.get("MSTK_BENCH_NOSAMPLER"):
            sampler.start()
        barrier()
        ms = timed(step_device, args.steps)
        barrier()
        ms_serial = timed(step_e2e, args.steps)
        barrier()
        e2e_runs = [loop_e2e(args.steps) for _ in range(3)]      # the loop is host-paced (one sample in three carries a 20-80 ms host pause): three runs of K calls, the fastest reported, all printed
        ms_e2e = min(e2e_runs)
        barrier()
        clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([ms, ms_e2e, ms_serial], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_serial = t.tolist()
    train = train_weak = parity = None
    if not args.no_train:
        del model
        torch.cuda.empty_cache()
        train = train_leg(args, rank, world, dev, barrier)
        if world > 1:
            # supplementary: the same step with the PER-GPU batch held at 256 (weak scaling of the training step)
            train_weak = train_leg(args, rank, world, dev, barrier, global_batch=TRAIN_GLOBAL_BATCH * world, rooflines=False)
            parity = ddp_parity(args, rank, world, dev)
        model = M.LDM(latent_dim=32, pretrained_path=None)          # fresh sampler model for the roofline leg
        model.load_state_dict(sd, strict=True)
        model.to(dev).eval().set_precision(args.precision)
    # config 5 (mel front end + transfer at 128x384) on every rank: batch-sharded like config 3, no collective
    mel = None
    if not args.no_mel:
        peaks_all = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks_all = json.load(f)
        except OSError:
            pass
        barrier()
        mel = mel_leg(model, dev, peaks_all, rank, cpu_baseline=(rank == 0 and world == 1 and not args.no_cpu_baseline))
        tm = torch.tensor([mel["mel"]["us"], mel["end_to_end"]["ms_per_call"]], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        us_mel, ms_c5 = tm.tolist()
        mel["aggregate"] = {"n_gpus": world, "clips": MEL_CLIPS * world, "mel_clips_per_s": MEL_CLIPS * world / (us_mel * 1e-6),
                            "config5_unet_denoise_steps_per_s": (NSTEPS - 1) * world / (ms_c5 * 1e-3),
                            "config5_clips_per_s": MEL_CLIPS * world / (ms_c5 * 1e-3), "timing": "max over ranks"}
    unet_steps = (NSTEPS - 1) * args.steps * world
    value = unet_steps / (ms * 1e-3)
    ms_pipelined = ms_e2e
    ms_e2e = min(ms_pipelined, ms_serial)                    # the pipelined loop is host-paced: on a noisy host the per-call figure can be the better one
    e2e_value = unet_steps / (ms_e2e * 1e-3)

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except OSError:
            pass
        with torch.no_grad():
            roof, rows = kernel_roofline(model, peaks)
        # single-GPU measurements (config 2 and the autoencoder stage do not shard): skipped in the N > 1 scaling runs
        vae = None if (args.no_vae or world > 1) else vae_leg(sd, dev, peaks, args.precision)
        vae32 = None if (args.no_vae or world > 1 or args.precision == "fp32") else vae_leg(sd, dev, peaks, "fp32", reps=5)
        ae = None if (args.no_train or world > 1) else ae_leg(sd, dev, args.precision)
        fp32 = None if (world > 1 or args.precision == "fp32" or args.no_fp32) else fp32_leg(sd, dev, content_d, style_d)
        sweep = None
        if world == 1 and not args.no_fp32 and roof and roof.get("flop_per_launch_avg"):
            sweep = batch_sweep(model, dev, peaks, roof["flop_per_launch_avg"] * roof["launches_per_unet_step"] / BATCH)
        base = None if (world > 1 or args.no_gpu_baseline) else gpu_baseline(sd, dev)
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            v, threads, dt = cpu_transfer_timing(10)
            cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": f"10 full config-3 calls (batch 64, 49 UNet+DDIM steps + VAE/style passes) of the oracle port, fp32, {dt:.1f} s"}
        unet_step_us = 1e3 * ms / args.steps / (NSTEPS - 1)

        def pick(d, *path):
            for p in path:
                if not isinstance(d, dict) or p not in d:
                    return None
                d = d[p]
            return d

        def r3(v):
            return None if v is None else float(f"{v:.4g}")

        def best(d, ks):
            vs = [v for v in (pick(d, k, "value") for k in ks) if v]
            return max(vs) if vs else None

        c3k = ("config3_eager_fp32_tf32", "config3_autocast_bf16", "config3_autocast_bf16_cudagraph")
        c4k = ("config4_eager_fp32_tf32", "config4_autocast_bf16")
        tv = pick(train, "value")
        # compact block, LAST in the line (the driver stores tails): everything BASELINE's metric names, with the bar beside it
        summary = {
            "sampling": {"steps_per_s": r3(value), "e2e_steps_per_s": r3(e2e_value), "unet_step_us": r3(unet_step_us),
                         "sum_isolated_kernel_us": r3(roof.get("sum_isolated_us_all_kernels")), "pdl_overlap_frac": r3(roof.get("pdl_overlap_frac")),
                         "launches_per_unet_step": roof.get("launches_per_unet_step_all_kernels"), "roofline_frac": r3(roof["frac"]),
                         "fp32_mode_steps_per_s": r3(pick(fp32, "value")),
                         "batch_sweep_frac_of_bf16_peak": None if not sweep else {str(r["batch"]): r3(r["frac_of_bf16_peak"]) for r in sweep},
                         "batch_sweep_unet_step_us": None if not sweep else {str(r["batch"]): r3(r["unet_step_us"]) for r in sweep}},
            "train_strong": {"global_batch": pick(train, "global_batch"), "samples_per_s": r3(tv), "ms_per_step": r3(pick(train, "ms_per_step")),
                             "e2e_samples_per_s": r3(pick(train, "e2e", "value")), "vggish_conv_roofline_frac": r3(pick(train, "roofline", "frac")),
                             "algorithmic_frac_of_bf16_peak": r3(19.6e9 * tv / 1e12 / (world * peaks.get("bf16_tflops", 1590.0))) if tv else None},
            "train_weak": {"global_batch": pick(train_weak, "global_batch"), "samples_per_s": r3(pick(train_weak, "value")),
                           "ms_per_step": r3(pick(train_weak, "ms_per_step")), "e2e_samples_per_s": r3(pick(train_weak, "e2e", "value"))},
            "ddp_parity": parity,
            "vae": {"eval_us": r3(pick(vae, "eval", "us_per_round_trip")), "eval_frac": r3(pick(vae, "eval", "frac")),
                    "train_us": r3(pick(vae, "train", "us_per_round_trip")), "train_frac": r3(pick(vae, "train", "frac")),
                    "fp32_eval_us": r3(pick(vae32, "eval", "us_per_round_trip")), "fp32_eval_frac": r3(pick(vae32, "eval", "frac")),
                    "fp32_train_us": r3(pick(vae32, "train", "us_per_round_trip")), "fp32_train_frac": r3(pick(vae32, "train", "frac"))},
            "config5": {"mel_us_64_clips": r3(pick(mel, "mel", "us")), "mel_hbm_frac": r3(pick(mel, "mel", "roofline", "frac")),
                        "mel_clips_per_s": r3(pick(mel, "aggregate", "mel_clips_per_s")),
                        "unet_denoise_steps_per_s": r3(pick(mel, "aggregate", "config5_unet_denoise_steps_per_s")),
                        "clips_per_s": r3(pick(mel, "aggregate", "config5_clips_per_s")),
                        "mel_cpu_clips_per_s": r3(pick(mel, "mel", "cpu_baseline", "clips_per_s")),
                        "inverse_clips_per_s": r3(pick(mel, "inverse", "clips_per_s")), "griffinlim_us_per_iteration": r3(pick(mel, "inverse", "us_per_iteration")),
                        "griffinlim_hbm_frac": r3(pick(mel, "inverse", "roofline", "frac")),
                        "inverse_cpu_clips_per_s": r3(pick(mel, "inverse", "cpu_baseline", "clips_per_s"))},
            "autoencoder_samples_per_s": r3(pick(ae, "value")),
            "gpu_baseline": None if base is None else {
                "what": "stock PyTorch (cuDNN/cuBLAS) on this B200, same inputs: eager fp32+TF32 / autocast bf16 / autocast bf16 in one CUDA graph",
                "config3_steps_per_s": [r3(pick(base, k, "value")) for k in c3k],
                "config3_speedup_vs_best": r3(value / best(base, c3k)) if best(base, c3k) else None,
                "config4_samples_per_s": [r3(pick(base, k, "value")) for k in c4k],
                "config4_speedup_vs_best": r3(tv / best(base, c4k)) if (tv and best(base, c4k)) else None,
                "config2_eval_us": [r3(pick(base, k, "us_per_round_trip")) for k in ("config2_eval_eager_fp32_tf32", "config2_eval_autocast_bf16")]},
            "cpu_steps_per_s": r3(pick(cpu, "value")), "n_gpus": world,
        }
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": workload_config({"precision": args.precision, "sharding": f"batch-sharded x{world}, no collective on the data path"}),
            "layers": rows,
            "train": train, "train_weak": train_weak, "vae": vae, "vae_fp32": vae32, "train_autoencoder": ae, "config5": mel, "fp32_mode": fp32, "batch_sweep": sweep,
            "gpu_baseline": base,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * content_h.numel() * 4,
                    "d2h_bytes_per_step": 2 * out_h[0].numel() * 4, "ms_per_step": ms_e2e / args.steps,
                    "how": "LDM.transfer_batches: pinned host batches in, pinned host images out; upload of call i+1 and downloads of "
                           "earlier calls on copy streams under call i (results handed out 6 calls late, ring of 8 pinned pairs); one event pair around all K calls, the L2 flush between calls inside it; "
                           "fastest of three runs of K calls (all in runs_ms_per_step), cyclic GC held off during the loop",
                    "method": "pipelined (transfer_batches)" if ms_pipelined <= ms_serial else "serial (one wrapper call at a time, event pair per call)",
                    "pipelined_value": unet_steps / (ms_pipelined * 1e-3), "runs_ms_per_step": [r / args.steps for r in e2e_runs],
                    "serial_value": unet_steps / (ms_serial * 1e-3), "serial_ms_per_step": ms_serial / args.steps},
            "gpu_launches": int(launches_per_step * args.steps * 2),
            "roofline": roof, "cpu_baseline": cpu, "clocks": clocks,
            "samples_steps_per_s": value * BATCH,
            "unet_step_us": unet_step_us,
            "summary": summary,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity["pass"]:
        raise SystemExit(f"bench.py: data-parallel self-check failed: {parity}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="mstk", choices=["mstk", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the training-step leg")
    ap.add_argument("--no-vae", action="store_true", help="skip the config-2 VAE round-trip leg")
    ap.add_argument("--no-mel", action="store_true", help="skip the config-5 mel front end + transfer leg")
    ap.add_argument("--no-fp32", action="store_true", help="skip the fp32-mode sampling leg")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the stock-PyTorch-on-this-GPU baseline legs")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps is None:
            args.steps = 10
        args.steps = min(args.steps, 20)     # each step (one full config-3 call) is ~1 s on 16 host cores: 10-20 s of CPU work
        run_reference(args)
    else:
        if args.steps is None:
            args.steps = 20
        run_cuda(args)


if __name__ == "__main__":
    main()
