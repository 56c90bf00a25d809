"""Drop-in surface of the reference's models/train.py training step on libmstk kernels.

    LDMTrainer(model, train_loader, device, lr=1e-4, style_loss_weight=0.1)      train.py:141-161
        .train_step(content_spec, style_spec) -> dict of 4 floats                 train.py:163-208
        .train_step_device(content, style, t=None, noise=None) -> dict of 0-dim CUDA tensors (no host sync)

The reference gets its gradients from autograd and its update from torch.optim.Adam under fp16
autocast + GradScaler.  Here forward, backward and update are one explicit stream of C-ABI kernel
launches (no autograd tape, no torch compute op on the path):

  * forward = LDM.forward's kernels with the intermediates kept (conv inputs, pre-add ReLU outputs,
    BatchNorm batch statistics, attention q / kv, the time MLP's e / pre / h);
  * gradient reaches the decoder, the UNet and the style encoder through MSE(recon, x) and
    MSE(eps_hat, eps) only: the encoder is frozen (model.py:340-341), the two VGGish terms are
    evaluated under torch.no_grad() in the reference (loss.py:78) and the KL term sees only the
    frozen encoder's output -- so those three are forward-only here too;
  * data gradients of the contractions run on the forward kernels with re-packed weights, weight
    gradients accumulate in a flat fp32 gradient arena in the parameters' own layouts, and the
    update is ONE fused Adam kernel over the flat parameter / gradient / moment arenas;
  * bf16 mode needs no loss scaling, so GradScaler is omitted (documented deviation; fp32 master
    weights and fp32 weight gradients as under autocast).

Data parallelism (one process per GPU): pass ``process_group`` / ``world_size``.  The batch is sharded;
the gradient arena is all-reduced in buckets (NCCL, side stream) as soon as the backward sweep has
finished the layers of a bucket, and the BatchNorm forward/backward statistics are all-reduced so every
rank normalises with the GLOBAL batch statistics (single-process parity).
        .train_epoch(epoch) -> (avg total, avg compression, avg denoising, avg style)   train.py:209-243
        .train(num_epochs)                                                         train.py:245-293
The epoch loop keeps the reference's contract (loader element layout, the ``training_iteration_noise``
factor in the averages, ReduceLROnPlateau(factor 0.5, patience 10) on the epoch loss, a checkpoint every
100 epochs) but accumulates the losses on the device -- one host read per epoch instead of one per step.
Plotting (matplotlib) and tqdm are outside the hot path.
"""
from __future__ import annotations

import os

from typing import Dict, List, Optional

import torch
import torch.nn as nn

from . import loss as L
from .config import config
from . import model as Mo
from . import ops
from .ops import ACT_NONE, ACT_RELU, ACT_TANH, MstkError, device_prefetch


class _Arena:
    """Flat fp32 storage for the trainable parameters (in backward-completion order), their gradients and
    the two Adam moments.  ``param.data`` / ``param.grad`` become views, so ``state_dict`` is unchanged."""

    def __init__(self, params: List[nn.Parameter], device):
        self.params = params
        self.offsets, off = [], 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() + 3) // 4 * 4             # keep every view 16-byte aligned
        self.size = off
        self.p = torch.zeros(off, dtype=torch.float32, device=device)
        self.g = torch.zeros(off, dtype=torch.float32, device=device)
        self.m = torch.zeros(off, dtype=torch.float32, device=device)
        self.v = torch.zeros(off, dtype=torch.float32, device=device)
        for p, o in zip(params, self.offsets):
            view = self.p[o:o + p.numel()].view(p.shape)
            view.copy_(p.data)
            p.data = view
            p.grad = self.g[o:o + p.numel()].view(p.shape)

    def span(self, first: nn.Parameter, last: nn.Parameter):
        i0 = next(i for i, p in enumerate(self.params) if p is first)
        i1 = next(i for i, p in enumerate(self.params) if p is last)
        return self.offsets[i0], self.offsets[i1] + (self.params[i1].numel() + 3) // 4 * 4


def _backward_order(model) -> List[nn.Parameter]:
    """trainable parameters in the order the backward sweep finishes them (decoder, UNet top-down, style encoder)"""
    d, u, s = model.decoder.decoder, model.unet, model.style_encoder
    mods = [d[6], d[4], d[3], d[1], d[0], u.dec1, u.dec2, u.dec3, u.dec4, u.bottleneck,
            u.cross_attention1.multihead_attn, u.enc4, u.cross_attention2.multihead_attn, u.enc3, u.enc2, u.time_mlp[3],
            u.time_mlp[1], u.enc1, s.enc6, s.enc5, s.enc4, s.enc3, s.enc2, s.enc1]
    out, seen = [], set()
    for m in mods:
        for p in m.parameters():
            if id(p) not in seen:
                seen.add(id(p))
                out.append(p)
    return out


class _PackRefresher:
    """Keeps the packed copies of the trainable weights valid across raw-pointer optimizer steps with ONE launch per step.

    The step's first run records every un-scaled pack (ops.PACK_RECORDER); afterwards, right behind the fused Adam / AdamW
    launch, `refresh()` re-packs all recorded copies of arena parameters in place (ops.PackPlan -> mstk_pack_multi) and drops
    only what it cannot refresh (BatchNorm folds, composed attention weights, captured sampling graphs), instead of
    dropping every cache entry and paying one pack launch in front of every convolution of the next step (52 of the 145
    launches of the config-4 step, and a weight prefetch that raced the pack kernel in front of it)."""

    def __init__(self, roots, arena_p: torch.Tensor, skip=()):
        self.roots, self.skip = list(roots), tuple(skip)
        self.lo = arena_p.data_ptr()
        self.hi = self.lo + arena_p.numel() * arena_p.element_size()
        self.entries, self.plan, self.keep_ids = [], None, set()
        self._saved, self._active, self._keepalive = None, False, []

    def begin(self) -> None:
        if self._active:                       # a step that raised between begin() and refresh()
            ops.PACK_RECORDER = self._saved
        self._saved, self._active = ops.PACK_RECORDER, True
        ops.PACK_RECORDER = []

    def _modules(self):
        for root in self.roots:
            for m in root.modules():
                if not any(m is k for k in self.skip):
                    yield m

    def refresh(self) -> None:
        new, ops.PACK_RECORDER, self._active = ops.PACK_RECORDER or [], self._saved, False
        if new and not torch.cuda.is_current_stream_capturing():
            cached = set()
            for m in self._modules():
                for _, val in m.__dict__.get("_mstk_cache", {}).values():
                    if isinstance(val, torch.Tensor):
                        cached.add(id(val))
            new = [e for e in new if id(e["dst"]) in cached]          # packs of skipped (frozen, kept) modules are not ours
            self.keep_ids |= {id(e["dst"]) for e in new}
            self._keepalive += [e["dst"] for e in new]
            mine = [e for e in new if self.lo <= e["src"].data_ptr() < self.hi]   # frozen weights need no refresh, only keeping
            if mine:
                self.entries += mine
                self.plan = ops.PackPlan(self.entries)
        if self.plan is not None:
            self.plan.run()
            for e in self.entries:             # fp32 mode hangs the hi/lo split of a packed copy on the copy itself (ops._conv_fp32_tc)
                if hasattr(e["dst"], "_mstk_w3"):
                    del e["dst"]._mstk_w3
        self.drop_stale()

    def drop_stale(self) -> None:
        """everything derived from the parameters that the plan does not refresh"""
        for m in self._modules():
            m.__dict__.pop("_graphs", None)
            cache = m.__dict__.get("_mstk_cache")
            if cache:
                for key in [k for k, (_, val) in cache.items() if not (isinstance(val, torch.Tensor) and id(val) in self.keep_ids)]:
                    del cache[key]


class PeerExchange:
    """The small fp64 all-reduces of the data-parallel step (BatchNorm statistics forward and backward) over peer-mapped
    memory: one single-CTA launch per rank instead of an NCCL collective (csrc/peer.cu).  One process per GPU; the exchange
    buffers are shared as CUDA IPC handles gathered over the process group.  `begin_step()` restarts the slot sequence: every
    rank must issue the same sequence of all_reduce_() calls per step."""

    MAXN, SLOTS = 512, 32

    def __init__(self, group, world_size: int, device):
        import ctypes as C
        import torch.distributed as dist
        from . import _lib
        self.lib = _lib.load()
        self.world, self.rank = int(world_size), dist.get_rank(group)
        if self.world > 8:
            raise MstkError("PeerExchange: at most 8 ranks (one NVSwitch domain)")
        own, handle = C.c_void_p(), (C.c_char * 64)()
        ops.check(self.lib.mstk_peer_alloc(C.byref(own), handle), "mstk_peer_alloc")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        self.own, self.mapped = own, []
        self.ctx = _lib.PeerCtx()
        self.ctx.rank, self.ctx.world = self.rank, self.world
        for r in range(self.world):
            if r == self.rank:
                self.ctx.bufs[r] = own.value
                continue
            p = C.c_void_p()
            ops.check(self.lib.mstk_peer_open(handles[r], C.byref(p)), "mstk_peer_open")
            self.mapped.append(p)
            self.ctx.bufs[r] = p.value
        self.slot = 0
        dist.barrier(group=group)                      # every rank has mapped every buffer before the first push

    def begin_step(self) -> None:
        self.slot = 0

    def fits(self, t: torch.Tensor) -> bool:
        return t.is_cuda and t.dtype == torch.float64 and t.is_contiguous() and 0 < t.numel() <= self.MAXN and self.slot < self.SLOTS

    def all_reduce_(self, t: torch.Tensor) -> None:
        import ctypes as C
        ops.check(self.lib.mstk_peer_allreduce_f64(t.data_ptr(), t.numel(), self.slot, C.byref(self.ctx), ops._stream()), "mstk_peer_allreduce_f64")
        self.slot += 1
        ops._count(1)

    def close(self) -> None:
        for p in self.mapped:
            self.lib.mstk_peer_close(p)
        self.mapped = []
        if self.own is not None:
            self.lib.mstk_peer_free(self.own)
            self.own = None


class GradReducer:
    """Data-parallel exchange of the training step: bucketed all-reduce (sum) of the flat gradient arena, issued per
    bucket behind the backward kernels that fill it and overlapped with the rest of the sweep on a side stream, plus
    the small all-reduces of BatchNorm statistics.  Works on CPU tensors too (gloo, synchronous) for the host-logic tests."""

    def __init__(self, grads: torch.Tensor, buckets, group=None, world_size: int = 1, peer: Optional["PeerExchange"] = None):
        self.g, self.buckets, self.group, self.world = grads, list(buckets), group, int(world_size)
        self.peer = peer
        self.stream = torch.cuda.Stream(device=grads.device) if (self.world > 1 and grads.is_cuda) else None
        lo = 0
        for a, b in self.buckets:
            if a != lo or b <= a:
                raise MstkError(f"GradReducer: buckets must tile the arena contiguously, got {self.buckets}")
            lo = b
        if lo != grads.numel():
            raise MstkError("GradReducer: buckets do not cover the arena")

    def all_reduce_small(self, t: torch.Tensor) -> None:
        if self.world > 1:
            if self.peer is not None and self.peer.fits(t):
                self.peer.all_reduce_(t)              # one single-CTA launch over peer-mapped memory (csrc/peer.cu)
                return
            import torch.distributed as dist
            dist.all_reduce(t, group=self.group)

    def reduce_bucket(self, i: int) -> None:
        if self.world <= 1 or i >= len(self.buckets):
            return
        import torch.distributed as dist
        lo, hi = self.buckets[i]
        if self.stream is None:
            dist.all_reduce(self.g[lo:hi], group=self.group)
            return
        ev = torch.cuda.Event()
        ev.record()
        self.stream.wait_event(ev)
        with torch.cuda.stream(self.stream):
            dist.all_reduce(self.g[lo:hi], group=self.group)

    def wait(self) -> None:
        if self.stream is not None:
            torch.cuda.current_stream().wait_stream(self.stream)


def bucket_spans(model, arena: "_Arena", num_buckets: int = 3):
    """contiguous [lo, hi) slices of the arena in backward-completion order: decoder + UNet decoder half,
    UNet bottleneck/attention/encoder half + time MLP, style encoder"""
    u, s, d = model.unet, model.style_encoder, model.decoder.decoder
    cuts = [(d[6].weight, u.dec4.bias), (u.bottleneck.weight, u.enc1.bias), (s.enc6.weight, s.enc1.bias)]
    spans = [arena.span(a, b) for a, b in cuts]
    n = max(1, min(3, num_buckets))
    merged = spans[:n - 1] + [(spans[n - 1][0], arena.size)]
    return merged


class PlateauScheduler:
    """torch.optim.lr_scheduler.ReduceLROnPlateau(mode='min', threshold=1e-4 'rel', cooldown=0, eps=1e-8) restated on a
    plain ``.lr`` attribute (the trainers here own a flat arena and one fused Adam launch, not a torch optimizer):
    train.py:159-161 (factor 0.5, patience 10, min_lr 0) and train.py:48 (config-driven, min_lr 1e-6)."""

    def __init__(self, owner, mode_factor: float = 0.5, patience: int = 10, min_lr: float = 0.0, threshold: float = 1e-4,
                 eps: float = 1e-8):
        if mode_factor >= 1.0:
            raise ValueError("Factor should be < 1.0.")
        self.owner, self.factor, self.patience, self.min_lr = owner, float(mode_factor), int(patience), float(min_lr)
        self.threshold, self.eps = float(threshold), float(eps)
        self.best, self.num_bad_epochs, self.last_epoch = float("inf"), 0, 0

    def step(self, metric) -> None:
        current = float(metric)
        self.last_epoch += 1
        if current < self.best * (1.0 - self.threshold):
            self.best, self.num_bad_epochs = current, 0
        else:
            self.num_bad_epochs += 1
        if self.num_bad_epochs > self.patience:
            new_lr = max(self.owner.lr * self.factor, self.min_lr)
            if self.owner.lr - new_lr > self.eps:
                self.owner.lr = new_lr
            self.num_bad_epochs = 0


class _SideStream:
    """mixin of the trainers"""

    # -- side stream for the parameter-gradient kernels -----------------------------------------------------------
    # Weight gradients and bias sums are leaves of the reverse sweep: nothing downstream reads them before the bucket
    # all-reduce / the optimizer.  On their own stream they run beside the data-gradient chain (the UNet's kernels are
    # 30-130 CTAs on 148 SMs); inside the captured step that is a parallel branch of the graph.
    def _side_begin(self):
        side = getattr(self, "_side", None)
        if side is not None:
            side.wait_stream(torch.cuda.current_stream())
        return side

    def _side_join(self) -> None:
        side = getattr(self, "_side", None)
        if side is not None:
            torch.cuda.current_stream().wait_stream(side)

    def _wgrad(self, x, g, dw, **kw) -> None:
        side = self._side_begin()
        if side is None:
            ops.wgrad(x, g, dw, **kw)
            return
        with torch.cuda.stream(side):
            ops.wgrad(x, g, dw, **kw)
        x.record_stream(side); g.record_stream(side)

    def _dbias(self, g, dbias) -> None:
        side = self._side_begin()
        if side is None:
            ops.act_bwd(g, None, act=ACT_NONE, dbias=dbias, want_g=False)
            return
        with torch.cuda.stream(side):
            ops.act_bwd(g, None, act=ACT_NONE, dbias=dbias, want_g=False)
        g.record_stream(side)


class LDMTrainer(_SideStream):
    def __init__(self, model, train_loader, device, lr=1e-4, style_loss_weight=0.1, process_group=None, world_size: int = 1,
                 num_buckets: int = 3):
        self.model = model.to(device)
        self.train_loader = train_loader
        self.device = torch.device(device)
        self.style_loss_weight = style_loss_weight
        self._lr = float(lr)
        self.scheduler = PlateauScheduler(self, mode_factor=0.5, patience=10)      # train.py:159-161
        self.checkpoint_dir = "models/pretrained"                                # train.py:266
        self.betas, self.eps = (0.9, 0.999), 1e-8          # torch.optim.Adam defaults (train.py:156)
        self.step_count = 0
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=device)
        # Whole-step CUDA graph: after `graph_warmup` eager steps the step (forward, losses, backward, bucketed all-reduce,
        # Adam) is captured once per input shape and replayed; t and the noise come from torch's graph-safe CUDA
        # generator exactly as in eager mode.  Only used when t / noise are not passed in.
        self.use_cuda_graph = True
        self.graph_warmup = 2
        self._graph = None
        self._eager_steps = 0
        self.group, self.world = process_group, int(world_size)
        self._gen = None
        if self.device.type != "cuda":
            raise MstkError("LDMTrainer: the training step runs on sm_100a kernels only (no CPU fallback)")
        order = _backward_order(self.model)
        trainable = [p for p in self.model.parameters() if p.requires_grad]
        if {id(p) for p in trainable} != {id(p) for p in order}:
            raise MstkError("LDMTrainer: expected exactly decoder + unet + style_encoder to be trainable (the reference freezes the "
                            "encoder, model.py:340-341, and VGGishFeatureLoss is frozen); call model._freeze_encoder() first")
        self.arena = _Arena(order, self.device)
        Mo.invalidate_packed(self.model)                      # packed-weight caches point at the old storages
        self.packs = _PackRefresher([self.model], self.arena.p, skip=(self.model.feature_loss_net,))
        # parameter-gradient kernels on a side stream (MSTK_WGRAD_STREAM=0: in line)
        self._side = torch.cuda.Stream(device=self.device) if os.environ.get("MSTK_WGRAD_STREAM", "1") != "0" else None
        # gradient buckets for the data-parallel all-reduce, in backward-completion order
        self.buckets = bucket_spans(self.model, self.arena, num_buckets)
        # BatchNorm statistic exchanges over peer-mapped memory instead of NCCL (MSTK_PEER_BN=0 keeps NCCL); falls back when the
        # ranks cannot map each other's memory (no peer access / IPC)
        self.peer = None
        if self.world > 1 and os.environ.get("MSTK_PEER_BN", "1") != "0":
            import torch.distributed as dist
            ok = torch.ones(1, dtype=torch.int32, device=self.device)
            try:
                self.peer = PeerExchange(self.group, self.world, self.device)
            except Exception as e:      # noqa: BLE001
                print(f"[mstk] peer-memory BatchNorm exchange unavailable ({e}); using NCCL", flush=True)
                ok.zero_()
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
            if int(ok) == 0 and self.peer is not None:
                self.peer.close()
                self.peer = None
        self.reducer = GradReducer(self.arena.g, self.buckets, self.group, self.world, peer=self.peer)
        if self.world > 1:
            import torch.distributed as dist
            dist.broadcast(self.arena.p, src=0, group=self.group)        # identical replicas
            for b in self.model.buffers():                                # BatchNorm running buffers
                if b.dtype.is_floating_point:
                    dist.broadcast(b, src=0, group=self.group)

    def close(self) -> None:
        """Drop the captured step.  A CUDA graph that holds NCCL kernels keeps the communicator busy:
        ``torch.distributed.destroy_process_group()`` blocks until every such graph is gone, so data-parallel callers
        close their trainers first."""
        self._graph = None
        self._eager_steps = 0
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        if getattr(self, "peer", None) is not None:
            self.reducer.peer = None
            self.peer.close()
            self.peer = None

    @property
    def lr(self) -> float:
        return self._lr

    @lr.setter
    def lr(self, value: float) -> None:
        """the learning rate is a launch parameter of the Adam kernel: a captured step is re-captured when it changes"""
        if float(value) != self._lr:
            self._lr = float(value)
            if self._graph is not None:
                self._graph = None
                self._eager_steps = self.graph_warmup

    # -- reference API ----------------------------------------------------------------------
    def train_epoch(self, epoch):
        """train.py:209-243: one pass over ``train_loader`` whose elements are ((content, label), (style, label));
        returns the four epoch averages (total, compression, denoising, style), each multiplied by
        config['training_iteration_noise'] as the reference does.  Losses are summed on the device."""
        self.model.train()
        num_batches = len(self.train_loader)
        acc = torch.zeros(4, dtype=torch.float64, device=self.device)
        for content_spec, style_spec in device_prefetch(self.train_loader, self.device, pick=lambda el: (el[0][0], el[1][0])):
            out = self.train_step_device(content_spec, style_spec)
            acc += torch.stack([out["total_loss"], out["compression_loss"], out["denoisinsg_loss"], out["style_loss"]]).double()
        k = config["training_iteration_noise"]
        return tuple(v / num_batches * k for v in acc.tolist())                  # ONE host sync per epoch

    def train(self, num_epochs):
        """train.py:245-293 without the matplotlib plots: epochs, ReduceLROnPlateau on the epoch's total loss,
        a state_dict checkpoint every 100 epochs into ``self.checkpoint_dir`` (reference: models/pretrained/;
        None skips the write).  Returns the per-epoch (total, compression, denoising, style) history."""
        history = []
        for epoch in range(num_epochs):
            losses = self.train_epoch(epoch)
            print(f"Epoch {epoch}: Train Loss = {losses[0]:.4f}")
            self.scheduler.step(losses[0])
            history.append(losses)
            print(f"Compression Loss: {losses[1]:.4f}")
            print(f"Denoisinsg Loss: {losses[2]:.4f}")
            print(f"Style Loss: {losses[3]:.4f}")
            rank0 = True
            if self.world > 1:
                import torch.distributed as dist
                rank0 = dist.get_rank(self.group) == 0                  # identical replicas: one writer
            if epoch % 100 == 0 and self.checkpoint_dir is not None and rank0:
                import os
                os.makedirs(self.checkpoint_dir, exist_ok=True)
                torch.save(self.model.state_dict(), os.path.join(self.checkpoint_dir, f"ldm_{epoch}.pth"))
        return history

    def train_step(self, content_spec, style_spec):
        out = self.train_step_device(content_spec, style_spec)
        vals = torch.stack([out[k] for k in ("compression_loss", "denoisinsg_loss", "style_loss", "total_loss")]).tolist()   # ONE host sync
        return dict(zip(("compression_loss", "denoisinsg_loss", "style_loss", "total_loss"), vals))

    def _G(self, param: torch.Tensor) -> torch.Tensor:
        """where the reverse sweep accumulates the gradient of ``param`` (the trainer: the arena view behind .grad)"""
        return param.grad

    def _sync_sums(self, sums):
        self.reducer.all_reduce_small(sums)

    def _reduce_bucket(self, stage: int):
        """the backward sweep finished stage 0 (decoder + UNet decoder half), 1 (rest of the UNet) or 2 (style encoder):
        launch the all-reduce of every bucket that is complete now (the last bucket always ends with stage 2)"""
        self._side_join()                       # the bucket's weight gradients run on the side stream
        n = len(self.buckets)
        for b in range(n):
            if (b if b < n - 1 else 2) == stage:
                self.reducer.reduce_bucket(b)

    # -- helpers ------------------------------------------------------------------------------
    @staticmethod
    def _dgrad(mod, name, conv, dy, *, w=None, out_dtype=None):
        """data gradient of `conv` (Conv2d / ConvTranspose2d / Linear) w.r.t. its input, on the forward kernels"""
        w = conv.weight if w is None else w
        dt = dy.dtype
        if isinstance(conv, nn.ConvTranspose2d):        # convT s2 -> conv s2, weight read as [co = cin_T][ci = cout_T]
            wp = mod._cached((name, "dg", dt), (conv.weight,), lambda: ops.pack_weight(w, dt))
            return ops.conv(dy, wp, None, ksize=conv.kernel_size[0], stride=2, out_dtype=out_dtype)
        if isinstance(conv, nn.Conv2d) and conv.stride[0] == 2:   # conv s2 -> convT s2, weight read as [cin_T = co][cout_T = ci]
            wp = mod._cached((name, "dg", dt), (conv.weight,), lambda: ops.pack_weight(w, dt, transposed=True))
            return ops.conv(dy, wp, None, ksize=3, stride=2, transposed=True, out_dtype=out_dtype)
        k = 1 if w.dim() == 2 else w.shape[2]           # stride-1 conv / Linear: flipped-tap, transposed pack
        wp = mod._cached((name, "dg", dt), (conv.weight if hasattr(conv, "weight") else w,), lambda: ops.pack_weight_dgrad(w, dt))
        return ops.conv(dy, wp, None, ksize=k, stride=1, out_dtype=out_dtype)

    def _dec_layer(self, name, conv, bn, x, adt):
        dec = self.model.decoder
        c = bn.num_features
        stats = ops.zeros_f64(2 * c, x.device)
        v = ops.conv(x, dec._packed(name, conv.weight, adt, True), dec._f32(name + ".b", conv.bias), ksize=4, stride=2, transposed=True,
                     act=ACT_NONE, stats=stats, out_dtype=adt)
        Mo._sync_bn_stats(stats)
        mean, invstd = torch.empty(c, device=x.device), torch.empty(c, device=x.device)
        rows = v.numel() // c
        h = ops.bn_apply(v, stats, bn.weight.detach(), bn.bias.detach(), count=rows * self.world, eps=bn.eps,
                         momentum=bn.momentum if bn.momentum is not None else 0.1, act=ACT_RELU, out_dtype=adt,
                         running_mean=bn.running_mean, running_var=bn.running_var, save_mean=mean, save_invstd=invstd)
        bn.num_batches_tracked.add_(1)
        return v, h, mean, invstd

    def _attn_fwd(self, ca, feat, style):
        e, m, adt = ca.embed_dim, ca.multihead_attn, feat.dtype
        b, h, w, _ = feat.shape
        kv = ca._project_style_kv(style)
        wq = ca._cached(("wq", adt), (m.in_proj_weight,), lambda: ops.pack_weight(m.in_proj_weight.detach()[:e], adt))
        bq = ca._cached(("bq",), (m.in_proj_bias,), lambda: m.in_proj_bias.detach()[:e].float().contiguous())
        q = ops.conv(feat, wq, bq, ksize=1)
        kv2 = kv.view(b, kv.shape[1] * kv.shape[2], 2 * e)
        o = ops.attention(q.view(b, h * w, e), kv2[:, :, :e], kv2[:, :, e:], ca.num_heads)
        out = ops.conv(o.view(b, h, w, e), ca._packed("wo", m.out_proj.weight, adt), ca._f32("bo", m.out_proj.bias), ksize=1)
        return out, dict(feat=feat, style=style, q=q, kv=kv2, o=o)

    def _attn_bwd(self, ca, sv, d_out):
        """d_out: gradient w.r.t. the attention output [b,h,w,e] -> (d_feat, d_style); parameter grads accumulated"""
        e, m = ca.embed_dim, ca.multihead_attn
        b, h, w, _ = d_out.shape
        feat, style, q, kv, o = sv["feat"], sv["style"], sv["q"], sv["kv"], sv["o"]
        self._dbias(d_out, self._G(m.out_proj.bias))
        self._wgrad(o.view(b, h, w, e), d_out, self._G(m.out_proj.weight), ksize=1)
        d_o = self._dgrad(ca, "wo", m.out_proj, d_out)
        dq, dkv = ops.attention_bwd(q.view(b, h * w, e), kv[:, :, :e], kv[:, :, e:], d_o.view(b, h * w, e), ca.num_heads)
        dq4, dkv4 = dq.view(b, h, w, e), dkv.view(b, style.shape[1], style.shape[2], 2 * e)
        gw, gb = self._G(m.in_proj_weight), self._G(m.in_proj_bias)
        self._dbias(dq4, gb[:e])
        self._dbias(dkv4, gb[e:])
        self._wgrad(feat, dq4, gw[:e], ksize=1)
        self._wgrad(style, dkv4, gw[e:], ksize=1)
        wq_, wkv_ = m.in_proj_weight.detach()[:e], m.in_proj_weight.detach()[e:]
        d_feat = ops.conv(dq4, ca._cached(("wq", "dg", dq.dtype), (m.in_proj_weight,), lambda: ops.pack_weight_dgrad(wq_, dq.dtype)), None, ksize=1)
        d_style = ops.conv(dkv4, ca._cached(("wkv", "dg", dq.dtype), (m.in_proj_weight,), lambda: ops.pack_weight_dgrad(wkv_, dq.dtype)), None, ksize=1)
        return d_feat, d_style

    def _relu_conv_bwd(self, mod, name, conv, x_in, y_relu, dy, dy2=None, need_dx=True):
        """backward of y = ReLU(conv(x) + b): bias / weight gradients accumulated, returns dL/dx"""
        g = ops.act_bwd(dy, y_relu, act=ACT_RELU, dy2=dy2, dbias=self._G(conv.bias))
        tr = isinstance(conv, nn.ConvTranspose2d)
        self._wgrad(x_in, g, self._G(conv.weight), ksize=conv.kernel_size[0], stride=conv.stride[0], transposed=tr)
        return self._dgrad(mod, name, conv, g) if need_dx else None

    # -- the step ---------------------------------------------------------------------------
    def train_step_device(self, content_spec, style_spec, t: Optional[torch.Tensor] = None, noise: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        """Public device-side step: eager for the first ``graph_warmup`` steps, afterwards a CUDA-graph replay.  The
        captured step reads t and the noise from static buffers: they are drawn (or, when passed in, copied) outside
        the graph, so the replayed step is the same function of (content, style, t, noise) as the eager one."""
        if not self.use_cuda_graph or not content_spec.is_cuda or torch.cuda.is_current_stream_capturing():
            return self._step_impl(content_spec, style_spec, t, noise)
        key = (tuple(content_spec.shape), tuple(style_spec.shape), self.model._precision())
        if self._graph is not None and self._graph["key"] != key:
            self._graph = None
            self._eager_steps = 0
        if self._graph is None:
            if self._eager_steps < self.graph_warmup:
                self._eager_steps += 1
                return self._step_impl(content_spec, style_spec, t, noise)
            B, H, W = content_spec.shape[0], content_spec.shape[2], content_spec.shape[3]
            lat = self.model.encoder.encoder[6].out_channels
            ent = {"key": key, "x": content_spec.detach().float().contiguous().clone(),
                   "s": style_spec.detach().float().contiguous().clone(),
                   "t": torch.zeros((B,), dtype=torch.int64, device=self.device),
                   "noise": torch.zeros((B, lat, H // 8, W // 8), dtype=torch.float32, device=self.device)}
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0, c0 = ops.LAUNCHES, self.step_count
            with ops.graph_capture(g):
                ent["out"] = self._step_impl(ent["x"], ent["s"], ent["t"], ent["noise"])
            ent["launches"] = ops.LAUNCHES - n0
            ops.LAUNCHES, self.step_count = n0, c0          # capture enqueued nothing
            ent["graph"] = g
            self._graph = ent
        ent = self._graph
        ent["x"].copy_(content_spec)
        ent["s"].copy_(style_spec)
        if t is None or noise is None:
            t_d, noise_d = self._draw(*ent["noise"].shape)
            t = t_d if t is None else t
            noise = noise_d if noise is None else noise
        ent["t"].copy_(t)
        ent["noise"].copy_(noise)
        ent["graph"].replay()
        self.step_count += 1
        ops._count(ent["launches"])
        self.packs.drop_stale()                         # the replay refreshed the packed copies; folds / sampling graphs are stale
        return ent["out"]

    def _step_impl(self, content_spec, style_spec, t: Optional[torch.Tensor] = None, noise: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        # BatchNorm statistics are global-batch statistics INSIDE the step only (Mo.bn_sync_group): module forwards
        # outside a training step never all-reduce
        if self.peer is not None:
            self.peer.begin_step()
        if getattr(self, "_zpool", None) is None:
            self._zpool = ops.ZeroPool(self.device)
        saved, ops.ZERO_POOL = ops.ZERO_POOL, self._zpool
        self._zpool.begin()
        try:
            with Mo.bn_sync_group(self.group, self.world, reduce=self.reducer.all_reduce_small):
                return self._step_body(content_spec, style_spec, t, noise)
        finally:
            ops.ZERO_POOL = saved

    def _draw(self, B: int, c: int, h: int, w: int):
        """t ~ randint(0, T) (train.py:172) and eps ~ randn (model.py:110) for this rank's shard.  Single process: torch's
        default CUDA generator, as the reference.  Data parallel: a per-rank generator (seed = initial seed + 1 + rank), so
        that the shards of one global batch do not repeat the same timesteps / noise."""
        gen = None
        if self.world > 1:
            if self._gen is None:
                import torch.distributed as dist
                self._gen = torch.Generator(device=self.device)
                self._gen.manual_seed(torch.initial_seed() + 1 + dist.get_rank(self.group))
            gen = self._gen
        t = torch.randint(0, self.model.num_timesteps, (B,), device=self.device, generator=gen)
        noise = torch.randn((B, c, h, w), dtype=torch.float32, device=self.device, generator=gen)
        return t, noise

    def _step_body(self, content_spec, style_spec, t: Optional[torch.Tensor] = None, noise: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        """One optimisation step on this rank's shard.  ``t`` [B] int64 and ``noise`` [B,C,h,w] (NCHW, the reference's
        randn_like shape) may be passed in (parity tests, data-parallel sharding of a globally drawn batch); otherwise
        they are drawn as the reference does (train.py:172, model.py:110).  Returns the four losses as 0-dim tensors."""
        model = self.model
        u, dec, sty = model.unet, model.decoder, model.style_encoder
        d = dec.decoder
        if not d[1].training:
            raise MstkError("LDMTrainer: call model.train() first (train.py:212); the step back-propagates through train-mode BatchNorm")
        adt = Mo._act_dtype(model._precision())
        lp = adt == torch.bfloat16
        xs = model._spec_nhwc(content_spec, "train_step")
        ss = model._spec_nhwc(style_spec, "train_step")
        dev, B = xs.device, xs.shape[0]
        ab = model.noise_scheduler.alpha_bar_t
        self.arena.g.zero_()                                                     # optimizer.zero_grad()
        self.packs.begin()
        if t is None or noise is None:
            lat = model.encoder.encoder[6].out_channels
            t_d, noise_d = self._draw(B, lat, xs.shape[1] // 8, xs.shape[2] // 8)
            t = t_d if t is None else t
            noise = noise_d if noise is None else noise
        t = t.to(device=dev, dtype=torch.int64).contiguous()

        sv = self._forward_keep(xs, ss, t, noise)
        (xs, ss, z0, S, noise_l, z_t, zin, l1, l2, temb, te_e, te_pre, te_h, z1, r2, z2, z3, a3, sv2, z4, a4, sv1, bt, r_d4, d4, r_d3, d3,
         r_d2, d2, eps_hat, dec_in, v1, h1, mean1, inv1, v2, h2, mean2, inv2, rec, t, adt, ab) = sv

        # ================= losses (train.py:182-186) =================
        side = self._side_begin()                   # three small reductions: beside the feature stack, joined before the sums
        if side is not None:
            main = torch.cuda.current_stream()
            with torch.cuda.stream(side), ops.workspace_slot(1):
                den = ops.mse(eps_hat, noise_l).squeeze(0)
                mse_rec = ops.mse(rec, xs).squeeze(0)
                kl = ops.kl(z0).squeeze(0)
            for tt in (den, mse_rec, kl):
                tt.record_stream(main)
            for tt in (eps_hat, noise_l, rec, xs, z0):
                tt.record_stream(side)
        else:
            den = ops.mse(eps_hat, noise_l).squeeze(0)
            mse_rec = ops.mse(rec, xs).squeeze(0)
            kl = ops.kl(z0).squeeze(0)
        fl = model.feature_loss_net
        H, W = xs.shape[1], xs.shape[2]
        three = torch.empty((3 * B, H, W, 1), dtype=torch.float32, device=dev)
        three[:B].copy_(xs); three[B:2 * B].copy_(rec); three[2 * B:].copy_(ss)
        perc = torch.zeros(1, dtype=torch.float32, device=dev)
        styl = torch.zeros(1, dtype=torch.float32, device=dev)
        # one pass of the frozen stack for all three images; perc: compression_loss's perceptual term (loss.py:40),
        # styl: style_loss (loss.py:104-106); each tap = one conv + one fused streaming kernel
        fl._loss3(three, B, perc, styl)
        self._side_join()
        comp = mse_rec + 0.1 * perc.squeeze(0) + 0.01 * kl
        total = comp + den + self.style_loss_weight * styl.squeeze(0)

        # ================= backward: d/drec of mean((rec - x)^2), d/deps_hat of mean((eps_hat - eps)^2) =================
        g_rec = ops.mse_grad(rec, xs, 2.0 / rec.numel(), torch.float32)
        self._backward_sweep(sv, g_rec, den_scale=2.0 / eps_hat.numel())

        # ================= update (torch.optim.Adam, train.py:156, 200) =================
        self._side_join()
        self.reducer.wait()
        self.step_count += 1
        self.step_dev.add_(1)                           # device-side copy of the step count: the update is graph-replayable
        ops.adam_step(self.arena.p, self.arena.g, self.arena.m, self.arena.v, lr=self.lr, beta1=self.betas[0], beta2=self.betas[1],
                      eps=self.eps, step=self.step_dev, grad_scale=1.0 / self.world)
        self.packs.refresh()                            # parameters changed under the packed copies / captured sampling graphs
        losses = torch.stack([comp, den, styl.squeeze(0), total])
        if self.world > 1:                              # report the GLOBAL batch's losses (equal shards: mean of the ranks' means)
            self.reducer.all_reduce_small(losses)
            losses = losses / self.world
        return dict(zip(("compression_loss", "denoisinsg_loss", "style_loss", "total_loss"), losses.unbind(0)))

    # -- the two halves of the step, shared with the autograd face of LDM.forward (autograd.py) -------------------------
    def _forward_keep(self, xs, ss, t, noise):
        """LDM.forward (model.py:355-379) on this rank's shard with every intermediate the reverse sweep needs kept.
        xs, ss: [B,H,W,1] fp32; t: [B] int64; noise: [B,C,h,w] fp32 NCHW (the reference's randn_like shape)."""
        model = self.model
        u, dec, sty = model.unet, model.decoder, model.style_encoder
        d = dec.decoder
        adt = Mo._act_dtype(model._precision())
        lp = adt == torch.bfloat16
        dev = xs.device
        ab = model.noise_scheduler.alpha_bar_t
        # the style encoder and the time MLP do not depend on the content path: they run on the side stream beside the
        # encoder, q-sample and the first two UNet layers, and join where the time embedding is added
        l1, l2 = u.time_mlp[1], u.time_mlp[3]
        side = self._side_begin()
        if side is not None:
            main = torch.cuda.current_stream()
            with torch.cuda.stream(side), ops.workspace_slot(1):                 # own split-K scratch: the branches run concurrently
                S = sty._forward_nhwc(ss)                                        # six post-ReLU maps
                temb, te_e, te_pre, te_h = ops.time_embed_train(t.to(torch.float32), l1.weight.detach(), l1.bias.detach(), l2.weight.detach(), l2.bias.detach())
            for tt in list(S) + [temb, te_e, te_pre, te_h]:
                tt.record_stream(main)
            ss.record_stream(side)
        z0 = model.encoder._forward_nhwc(xs)                                     # frozen; BatchNorm in train mode (train.py:212)
        _, h, w, c = z0.shape
        if side is None:
            S = sty._forward_nhwc(ss)
        noise_l = ops.nchw_to_nhwc(noise.to(dev).float().contiguous(), torch.float32)
        z_t, z_t_lp = ops.q_sample(z0, noise_l, ab, t, lp=lp)
        zin = z_t_lp if lp else z_t
        if side is None:
            temb, te_e, te_pre, te_h = ops.time_embed_train(t.to(torch.float32), l1.weight.detach(), l1.bias.detach(), l2.weight.detach(), l2.bias.detach())
        z1 = u._conv("enc1", zin, act=ACT_RELU)
        r2 = u._conv("enc2", z1, act=ACT_RELU)
        self._side_join()
        z2 = ops.add_bcast(r2, add_nc=temb)
        z3 = u._conv("enc3", z2, act=ACT_RELU)
        a3, sv2 = self._attn_fwd(u.cross_attention2, z3, S[4])
        z4 = u._conv("enc4", a3, act=ACT_RELU)
        a4, sv1 = self._attn_fwd(u.cross_attention1, z4, S[5])
        bt = u._conv("bottleneck", a4, act=ACT_RELU)
        r_d4 = u._conv("dec4", bt, act=ACT_RELU); d4 = ops.add_bcast(r_d4, add=z3)
        r_d3 = u._conv("dec3", d4, act=ACT_RELU); d3 = ops.add_bcast(r_d3, add=z2)
        r_d2 = u._conv("dec2", d3, act=ACT_RELU); d2 = ops.add_bcast(r_d2, add=z1)
        eps_hat = u._conv("dec1", d2, act=ACT_NONE, out_dtype=torch.float32)
        z0p, z0p_lp = ops.predict_x0(z_t, eps_hat, ab, t, lp=lp)
        dec_in = z0p_lp if lp else z0p
        v1, h1, mean1, inv1 = self._dec_layer("d0", d[0], d[1], dec_in, adt)
        v2, h2, mean2, inv2 = self._dec_layer("d3", d[3], d[4], h1, adt)
        rec = ops.conv(h2, dec._packed("d6", d[6].weight, adt, True), dec._f32("d6.b", d[6].bias), ksize=4, stride=2, transposed=True,
                       act=ACT_TANH, out_dtype=torch.float32, out_scale=0.5, out_shift=0.5)      # reconstructed = (tanh + 1) / 2

        return (xs, ss, z0, S, noise_l, z_t, zin, l1, l2, temb, te_e, te_pre, te_h, z1, r2, z2, z3, a3, sv2, z4, a4, sv1, bt, r_d4, d4, r_d3, d3,
                r_d2, d2, eps_hat, dec_in, v1, h1, mean1, inv1, v2, h2, mean2, inv2, rec, t, adt, ab)

    def _backward_sweep(self, sv, g_rec, den_scale: float = 0.0, g_eps_direct: Optional[torch.Tensor] = None) -> None:
        """Reverse sweep from g_rec = dL/d(reconstructed) [B,H,W,1] fp32 and the gradient on eps_hat, which gathers
        den_scale * (eps_hat - eps) (the denoising loss), ``g_eps_direct`` (an external gradient on noise_pred, NHWC in
        the activation dtype) and the reconstruction path through x0_pred (model.py:117-124).  Parameter gradients are
        ACCUMULATED into self._G(param)."""
        model = self.model
        u, dec, sty = model.unet, model.decoder, model.style_encoder
        d = dec.decoder
        (xs, ss, z0, S, noise_l, z_t, zin, l1, l2, temb, te_e, te_pre, te_h, z1, r2, z2, z3, a3, sv2, z4, a4, sv1, bt, r_d4, d4, r_d3, d3,
         r_d2, d2, eps_hat, dec_in, v1, h1, mean1, inv1, v2, h2, mean2, inv2, rec, t, adt, ab) = sv
        # reconstruction path: d/drec of mean((rec - x)^2); tanh and the (y+1)/2 folded into act_bwd
        g6 = ops.act_bwd(g_rec, rec, act=ACT_TANH, dbias=self._G(d[6].bias), y_scale=0.5, y_shift=0.5)
        self._wgrad(h2, g6, self._G(d[6].weight), ksize=4, stride=2, transposed=True)
        w6 = dec._cached(("d6", "dg32"), (d[6].weight,), lambda: ops.pack_weight(d[6].weight, torch.float32))
        dh2 = ops.conv(g6, w6, None, ksize=4, stride=2, out_dtype=adt)
        dv2 = ops.bn_bwd(dh2, v2, mean2, inv2, d[4].weight.detach(), d[4].bias.detach(), act=ACT_RELU, dgamma=self._G(d[4].weight),
                         dbeta=self._G(d[4].bias), world=self.world, sync=self._sync_sums)
        self._dbias(dv2, self._G(d[3].bias))
        self._wgrad(h1, dv2, self._G(d[3].weight), ksize=4, stride=2, transposed=True)
        dh1 = self._dgrad(dec, "d3", d[3], dv2)
        dv1 = ops.bn_bwd(dh1, v1, mean1, inv1, d[1].weight.detach(), d[1].bias.detach(), act=ACT_RELU, dgamma=self._G(d[1].weight),
                         dbeta=self._G(d[1].bias), world=self.world, sync=self._sync_sums)
        self._dbias(dv1, self._G(d[0].bias))
        self._wgrad(dec_in, dv1, self._G(d[0].weight), ksize=4, stride=2, transposed=True)
        dz0p = self._dgrad(dec, "d0", d[0], dv1)
        # eps_hat gathers the denoising loss and the reconstruction path through x0_pred (model.py:117-124)
        g_eps = ops.eps_grad(dz0p, eps_hat, noise_l, ab, t, den_scale, adt)
        if g_eps_direct is not None:
            g_eps = ops.add_bcast(g_eps, add=g_eps_direct)
        # ---- UNet, top-down ----
        self._dbias(g_eps, self._G(u.dec1.bias))
        self._wgrad(d2, g_eps, self._G(u.dec1.weight), ksize=3)
        dd2 = self._dgrad(u, "dec1", u.dec1, g_eps)
        dd3 = self._relu_conv_bwd(u, "dec2", u.dec2, d3, r_d2, dd2)
        dd4 = self._relu_conv_bwd(u, "dec3", u.dec3, d4, r_d3, dd3)
        dbt = self._relu_conv_bwd(u, "dec4", u.dec4, bt, r_d4, dd4)
        self._reduce_bucket(0)
        da4 = self._relu_conv_bwd(u, "bottleneck", u.bottleneck, a4, bt, dbt)
        dz4, ds6_attn = self._attn_bwd(u.cross_attention1, sv1, da4)
        da3 = self._relu_conv_bwd(u, "enc4", u.enc4, a3, z4, dz4)
        dz3a, ds5_attn = self._attn_bwd(u.cross_attention2, sv2, da3)
        dz2c = self._relu_conv_bwd(u, "enc3", u.enc3, z2, z3, dz3a, dy2=dd4)        # + skip gradient of dec4's add
        # z2 = ReLU(enc2(z1)) + t_emb: its gradient is dz2c + dd3 (skip of dec3's add)
        dtemb = ops.sum_hw(dz2c, dd3)
        dz1c = self._relu_conv_bwd(u, "enc2", u.enc2, z1, r2, dz2c, dy2=dd3)
        self._relu_conv_bwd(u, "enc1", u.enc1, zin, z1, dz1c, dy2=dd2, need_dx=False)
        # time MLP (model.py:170-175): Linear2 <- GELU <- Linear1 <- sinusoid
        nb = dtemb.shape[0]
        dt4 = dtemb.view(nb, 1, 1, -1)
        self._dbias(dt4, self._G(l2.bias))
        self._wgrad(te_h.view(nb, 1, 1, -1), dt4, self._G(l2.weight), ksize=1)
        dh = self._dgrad(u, "t2", l2, dt4)
        dpre = ops.gelu_bwd(dh.view(nb, -1), te_pre).view(nb, 1, 1, -1)
        self._dbias(dpre, self._G(l1.bias))
        self._wgrad(te_e.view(nb, 1, 1, -1), dpre, self._G(l1.weight), ksize=1)
        self._reduce_bucket(1)
        # ---- style encoder: s6 feeds attention1, s5 feeds attention2 and enc6 ----
        ds5c = self._relu_conv_bwd(sty, "enc6", sty.enc6, S[4], S[5], ds6_attn)
        dsk = self._relu_conv_bwd(sty, "enc5", sty.enc5, S[3], S[4], ds5c, dy2=ds5_attn)
        dsk = self._relu_conv_bwd(sty, "enc4", sty.enc4, S[2], S[3], dsk)
        dsk = self._relu_conv_bwd(sty, "enc3", sty.enc3, S[1], S[2], dsk)
        dsk = self._relu_conv_bwd(sty, "enc2", sty.enc2, S[0], S[1], dsk)
        self._relu_conv_bwd(sty, "enc1", sty.enc1, ss, S[0], dsk, need_dx=False)
        self._reduce_bucket(2)


# =====================================================================================================
# Autoencoder stage: models/train.py:28-138
# =====================================================================================================
def _conv_bn_train(mod, name, conv, bn, x, *, transposed: bool, ksize: int, act: int, mid_dtype, out_dtype, first: bool = False):
    """conv / convT (stride 2) with the batch sums in its epilogue, then the normalise(+ReLU) pass; keeps what the
    backward sweep needs: the pre-BatchNorm map v and the batch mean / inverse std."""
    c = bn.num_features
    stats = ops.zeros_f64(2 * c, x.device)
    wp = mod._packed(name, conv.weight, torch.float32 if first else x.dtype, transposed)
    v = ops.conv(x, wp, mod._f32(name + ".b", conv.bias), ksize=ksize, stride=2, transposed=transposed, act=ACT_NONE, stats=stats,
                 out_dtype=mid_dtype)
    mean, invstd = torch.empty(c, device=x.device), torch.empty(c, device=x.device)
    h = ops.bn_apply(v, stats, bn.weight.detach(), bn.bias.detach(), count=v.numel() // c, eps=bn.eps,
                     momentum=bn.momentum if bn.momentum is not None else 0.1, act=act, out_dtype=out_dtype,
                     running_mean=bn.running_mean, running_var=bn.running_var, save_mean=mean, save_invstd=invstd)
    bn.num_batches_tracked.add_(1)
    return v, h, mean, invstd


class AutoencoderTrainer(_SideStream):
    """One optimisation step of the reference's autoencoder stage (train.py:47, 68-83) on libmstk kernels:
    latent = encoder(x); recon = decoder(latent) (tanh range, NOT rescaled -- train.py:71-75 compares it with the
    spectrogram directly); loss = compression_loss(x, recon, latent, feature_extractor); AdamW over encoder + decoder.

    Gradient reaches the parameters through MSE(recon, x) and through 0.01 * KL(latent) (encoder only); the feature
    stack of the perceptual term runs under torch.no_grad() in the reference (loss.py:78) and is forward-only here.
    Forward, backward and update are an explicit stream of C-ABI launches, as in LDMTrainer, replayed as one CUDA graph
    after two eager steps.  Single process."""

    def __init__(self, encoder, decoder, feature_extractor, device, lr=5e-4, weight_decay=1e-2):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise MstkError("AutoencoderTrainer: the training step runs on sm_100a kernels only (no CPU fallback)")
        self.encoder, self.decoder = encoder.to(device), decoder.to(device)
        self.feature_extractor = feature_extractor.to(device) if feature_extractor is not None else None
        self._lr, self.weight_decay = float(lr), float(weight_decay)       # torch.optim.AdamW defaults (train.py:47)
        self.betas, self.eps = (0.9, 0.999), 1e-8
        self.step_count = 0
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=device)
        # whole-step CUDA graph after `graph_warmup` eager steps, re-captured when the input shape, the precision or
        # the learning rate (a launch parameter of the AdamW kernel) changes
        self.use_cuda_graph, self.graph_warmup = True, 2
        self._graph, self._eager_steps = None, 0
        e, d = self.encoder.encoder, self.decoder.decoder
        # backward-completion order: decoder top-down, then the encoder top-down
        order = [d[6].weight, d[6].bias, d[4].weight, d[4].bias, d[3].weight, d[3].bias, d[1].weight, d[1].bias, d[0].weight, d[0].bias,
                 e[7].weight, e[7].bias, e[6].weight, e[6].bias, e[4].weight, e[4].bias, e[3].weight, e[3].bias, e[1].weight, e[1].bias,
                 e[0].weight, e[0].bias]
        if not all(p.requires_grad for p in order):
            raise MstkError("AutoencoderTrainer: every encoder / decoder parameter must require grad (train.py:47)")
        self.arena = _Arena(order, self.device)
        Mo.invalidate_packed(self.encoder); Mo.invalidate_packed(self.decoder)
        self.packs = _PackRefresher([self.encoder, self.decoder], self.arena.p)
        self._side = torch.cuda.Stream(device=self.device) if os.environ.get("MSTK_WGRAD_STREAM", "1") != "0" else None

    @property
    def lr(self) -> float:
        return self._lr

    @lr.setter
    def lr(self, value: float) -> None:
        if float(value) != self._lr:
            self._lr = float(value)
            if self._graph is not None:
                self._graph, self._eager_steps = None, self.graph_warmup

    def train_step(self, spectrogram) -> float:
        return float(self.train_step_device(spectrogram))

    def train_step_device(self, spectrogram) -> torch.Tensor:
        """eager for the first steps, afterwards a replay of the captured step; returns the loss as a 0-dim CUDA tensor"""
        if not self.use_cuda_graph or not spectrogram.is_cuda or torch.cuda.is_current_stream_capturing():
            return self._step_impl(spectrogram)
        key = (tuple(spectrogram.shape), self.encoder._precision(), self.feature_extractor is not None)
        if self._graph is not None and self._graph["key"] != key:
            self._graph, self._eager_steps = None, 0
        if self._graph is None:
            if self._eager_steps < self.graph_warmup:
                self._eager_steps += 1
                return self._step_impl(spectrogram)
            ent = {"key": key, "x": spectrogram.detach().float().contiguous().clone()}
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0, c0 = ops.LAUNCHES, self.step_count
            with ops.graph_capture(g):
                ent["out"] = self._step_impl(ent["x"])
            ent["launches"] = ops.LAUNCHES - n0
            ops.LAUNCHES, self.step_count = n0, c0          # capture enqueued nothing
            ent["graph"] = g
            self._graph = ent
        ent = self._graph
        ent["x"].copy_(spectrogram)
        ent["graph"].replay()
        self.step_count += 1
        ops._count(ent["launches"])
        self.packs.drop_stale()
        return ent["out"]

    def _step_impl(self, spectrogram) -> torch.Tensor:
        enc, dec = self.encoder, self.decoder
        e, d = enc.encoder, dec.decoder
        if not (e[1].training and d[1].training):
            raise MstkError("AutoencoderTrainer: call encoder.train() and decoder.train() first (train.py:60-61)")
        Mo._need_cuda(spectrogram, "AutoencoderTrainer")
        x = spectrogram.detach().float().contiguous()
        if x.dim() != 4 or x.shape[1] != 1 or x.shape[2] % 8 or x.shape[3] % 8:
            raise MstkError(f"AutoencoderTrainer: expected a [B,1,H,W] spectrogram with H, W multiples of 8, got {tuple(x.shape)}")
        B, _, H, W = x.shape
        xs = x.view(B, H, W, 1)
        adt = Mo._act_dtype(enc._precision())
        self.arena.g.zero_()                                                     # optimizer.zero_grad()
        self.packs.begin()

        # ---- forward (model.py:16-28, 37-49), intermediates kept ----
        v0, h0, m0, i0 = _conv_bn_train(enc, "c0", e[0], e[1], xs, transposed=False, ksize=3, act=ACT_RELU, mid_dtype=adt, out_dtype=adt, first=True)
        v1, h1, m1, i1 = _conv_bn_train(enc, "c3", e[3], e[4], h0, transposed=False, ksize=3, act=ACT_RELU, mid_dtype=adt, out_dtype=adt)
        v2, z, m2, i2 = _conv_bn_train(enc, "c6", e[6], e[7], h1, transposed=False, ksize=3, act=ACT_NONE, mid_dtype=torch.float32,
                                       out_dtype=torch.float32)
        dec_in = z if adt == torch.float32 else ops.cast(z, adt)
        u1, g1, n1, j1 = _conv_bn_train(dec, "d0", d[0], d[1], dec_in, transposed=True, ksize=4, act=ACT_RELU, mid_dtype=adt, out_dtype=adt)
        u2, g2, n2, j2 = _conv_bn_train(dec, "d3", d[3], d[4], g1, transposed=True, ksize=4, act=ACT_RELU, mid_dtype=adt, out_dtype=adt)
        rec = ops.conv(g2, dec._packed("d6", d[6].weight, adt, True), dec._f32("d6.b", d[6].bias), ksize=4, stride=2, transposed=True,
                       act=ACT_TANH, out_dtype=torch.float32)

        # ---- compression_loss (loss.py:34-45) ----
        loss = ops.mse(rec, xs).squeeze(0) + 0.01 * ops.kl(z).squeeze(0)
        if self.feature_extractor is not None:
            loss = loss + 0.1 * L.perceptual_loss(x, rec.view(B, 1, H, W), config["compression_feature_extractor"],
                                                  feature_extractor=self.feature_extractor)

        # ---- backward: decoder ----
        g_rec = ops.mse_grad(rec, xs, 2.0 / rec.numel(), torch.float32)
        g6 = ops.act_bwd(g_rec, rec, act=ACT_TANH, dbias=d[6].bias.grad)
        self._wgrad(g2, g6, d[6].weight.grad, ksize=4, stride=2, transposed=True)
        w6 = dec._cached(("d6", "dg32"), (d[6].weight,), lambda: ops.pack_weight(d[6].weight, torch.float32))
        dg2 = ops.conv(g6, w6, None, ksize=4, stride=2, out_dtype=adt)
        du2 = ops.bn_bwd(dg2, u2, n2, j2, d[4].weight.detach(), d[4].bias.detach(), act=ACT_RELU, dgamma=d[4].weight.grad, dbeta=d[4].bias.grad)
        self._dbias(du2, d[3].bias.grad)
        self._wgrad(g1, du2, d[3].weight.grad, ksize=4, stride=2, transposed=True)
        dg1 = LDMTrainer._dgrad(dec, "d3", d[3], du2)
        du1 = ops.bn_bwd(dg1, u1, n1, j1, d[1].weight.detach(), d[1].bias.detach(), act=ACT_RELU, dgamma=d[1].weight.grad, dbeta=d[1].bias.grad)
        self._dbias(du1, d[0].bias.grad)
        self._wgrad(dec_in, du1, d[0].weight.grad, ksize=4, stride=2, transposed=True)
        dz = LDMTrainer._dgrad(dec, "d0", d[0], du1)
        # ---- backward: encoder; the latent also carries 0.01 * KL ----
        gz = ops.kl_grad(z, dz, 0.01 / z.numel())
        dv2 = ops.bn_bwd(gz, v2, m2, i2, e[7].weight.detach(), e[7].bias.detach(), act=ACT_NONE, dgamma=e[7].weight.grad, dbeta=e[7].bias.grad,
                         dv_dtype=adt)
        self._dbias(dv2, e[6].bias.grad)
        self._wgrad(h1, dv2, e[6].weight.grad, ksize=3, stride=2)
        dh1 = LDMTrainer._dgrad(enc, "c6", e[6], dv2)
        dv1 = ops.bn_bwd(dh1, v1, m1, i1, e[4].weight.detach(), e[4].bias.detach(), act=ACT_RELU, dgamma=e[4].weight.grad, dbeta=e[4].bias.grad)
        self._dbias(dv1, e[3].bias.grad)
        self._wgrad(h0, dv1, e[3].weight.grad, ksize=3, stride=2)
        dh0 = LDMTrainer._dgrad(enc, "c3", e[3], dv1)
        dv0 = ops.bn_bwd(dh0, v0, m0, i0, e[1].weight.detach(), e[1].bias.detach(), act=ACT_RELU, dgamma=e[1].weight.grad, dbeta=e[1].bias.grad)
        self._dbias(dv0, e[0].bias.grad)
        self._wgrad(xs, dv0, e[0].weight.grad, ksize=3, stride=2)

        # ---- update (torch.optim.AdamW, train.py:47, 81-83) ----
        self._side_join()
        self.step_count += 1
        self.step_dev.add_(1)
        ops.adam_step(self.arena.p, self.arena.g, self.arena.m, self.arena.v, lr=self.lr, beta1=self.betas[0], beta2=self.betas[1],
                      eps=self.eps, step=self.step_dev, weight_decay=self.weight_decay)
        self.packs.refresh()                                      # parameters changed under the packed copies
        return loss


def train_autoencoder(config, train_loader=None, test_loader=None, checkpoint_dir="models/pretrained"):
    """models/train.py:28-138 without dataset preparation and plotting: ``train_loader`` / ``test_loader`` yield
    elements whose [0] is a [B,1,H,W] spectrogram batch (the reference builds them with prepare_dataset(config),
    which is dataset I/O outside the path).  AdamW + ReduceLROnPlateau on the validation loss, best-validation and
    final checkpoints of encoder.pth / decoder.pth.  Returns (encoder, decoder, train_losses, val_losses)."""
    import os
    if not torch.cuda.is_available():
        raise RuntimeError("GPU not available, please check your setup.")           # train.py:32-36
    if train_loader is None or test_loader is None:
        raise MstkError("train_autoencoder: pass train_loader and test_loader (dataset preparation is outside the hot path)")
    device = torch.device("cuda")
    encoder = Mo.SpectrogramEncoder(config["latent_dim_encoder"]).to(device)
    decoder = Mo.SpectrogramDecoder(config["latent_dim_encoder"]).to(device)
    feature_extractor = L.VGGishFeatureLoss().to(device)
    tr = AutoencoderTrainer(encoder, decoder, feature_extractor, device, lr=config["learning_rate"])
    scheduler = PlateauScheduler(tr, mode_factor=config["learning_rate_factor"], patience=config["learning_rate_patience"],
                                 min_lr=config["learning_rate_min"])

    def save():
        if checkpoint_dir is not None:
            os.makedirs(checkpoint_dir, exist_ok=True)
            torch.save(encoder.state_dict(), os.path.join(checkpoint_dir, "encoder.pth"))
            torch.save(decoder.state_dict(), os.path.join(checkpoint_dir, "decoder.pth"))

    train_losses, val_losses, best_val = [], [], float("inf")
    for epoch in range(config["num_epochs"]):
        encoder.train(); decoder.train()
        acc = torch.zeros((), dtype=torch.float64, device=device)
        for spec in device_prefetch(train_loader, device, pick=lambda el: el[0]):
            acc += tr.train_step_device(spec).double()
        train_losses.append(float(acc) / len(train_loader))
        encoder.eval(); decoder.eval()
        acc = torch.zeros((), dtype=torch.float64, device=device)
        with torch.no_grad():
            for spec in device_prefetch(test_loader, device, pick=lambda el: el[0]):
                latent = encoder(spec)
                acc += L.compression_loss(spec, decoder(latent), latent, feature_extractor).double()
        val_losses.append(float(acc) / len(test_loader))
        scheduler.step(val_losses[-1])
        if val_losses[-1] < best_val:
            best_val = val_losses[-1]
            save()
        print(f"Epoch: {epoch}")
        print(f"Average Train Loss: {train_losses[-1]:.6f}")
        print(f"Average Val Loss: {val_losses[-1]:.6f}")
        print(f"Learning Rate: {tr.lr:.6f}")
    save()
    return encoder, decoder, train_losses, val_losses
