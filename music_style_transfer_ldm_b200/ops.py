"""Thin functional layer over the C ABI: torch tensors are only device-memory handles.

Activations are NHWC (``[N, H, W, C]`` contiguous), fp32 or bf16.  Every function
enqueues on torch's current CUDA stream (so the calls are CUDA-graph capturable) and
raises ``MstkError`` on failure; nothing here computes with PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import (ACT_NONE, ACT_RELU, ACT_TANH, BACKEND_AUTO, BACKEND_SIMT, BACKEND_TCGEN05, BF16, F32,
                   MstkError, check)

TORCH_DTYPE = {F32: torch.float32, BF16: torch.bfloat16}


def dtype_code(t_or_dtype) -> int:
    d = t_or_dtype.dtype if isinstance(t_or_dtype, torch.Tensor) else t_or_dtype
    if d == torch.float32:
        return F32
    if d == torch.bfloat16:
        return BF16
    raise MstkError(f"unsupported dtype {d}")


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _dev(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise MstkError(f"{name}: expected a CUDA tensor (the kernels have no CPU path)")
    if not t.is_contiguous():
        raise MstkError(f"{name}: expected a contiguous tensor")
    return t


_workspaces = {}
WORKSPACE_BYTES = 32 << 20

# bookkeeping for bench.py: kernels launched through this module (graph replays are added by
# the caller that replays them) and, when RECORD is a list, a replayable record of conv calls
LAUNCHES = 0
RECORD = None
RECORD_NOTE = ""        # free-form tag copied into the records made while it is set (bench.py: "hoisted" = once per loop)


def _count(n: int = 1) -> None:
    global LAUNCHES
    LAUNCHES += n


def tc_eligible(dtype, cin: int, cout: int, stride: int, transposed: bool, h: int, w: int) -> bool:
    """mirror of conv_tc_supported() in csrc/conv_tc.cu (which kernel AUTO picks)"""
    if dtype != torch.bfloat16 or cin % 32 or cout % 16:
        return False
    if not transposed and stride == 2 and (h % 2 or w % 2):
        return False
    return True


def device_prefetch(loader, device, pick=lambda element: element):
    """Iterates ``loader`` one element ahead: yields ``pick(element)`` (a tensor or a tuple / list of tensors) on ``device``
    with the upload of the NEXT element already enqueued on a copy stream, so the host-to-device copy of step i+1 runs under
    the kernels of step i (33.5 MB per config-4 step: 0.6 ms of PCIe time beside a 6.9 ms step).  Pinned source tensors copy
    asynchronously; pageable ones work too (the copy then blocks the host, as ``.to(device)`` always does).  Tensors cross
    from the copy stream to the caller's stream through an event and ``record_stream``."""
    device = torch.device(device)
    if device.type != "cuda":
        for element in loader:
            item = pick(element)
            yield item.to(device) if isinstance(item, torch.Tensor) else tuple(t.to(device) for t in item)
        return
    copy = torch.cuda.Stream(device)

    def upload(element):
        item = pick(element)
        single = isinstance(item, torch.Tensor)
        with torch.cuda.stream(copy):
            moved = tuple(t.to(device, non_blocking=True) for t in ((item,) if single else item))
            done = torch.cuda.Event()
            done.record(copy)
        return moved, done, single

    it = iter(loader)
    try:
        ahead = upload(next(it))
    except StopIteration:
        return
    while ahead is not None:
        moved, done, single = ahead
        try:
            ahead = upload(next(it))              # enqueued before the consumer launches the step that uses `moved`
        except StopIteration:
            ahead = None
        main = torch.cuda.current_stream(device)
        main.wait_event(done)
        for t in moved:
            t.record_stream(main)
        yield moved[0] if single else moved


class graph_capture:
    """``torch.cuda.graph`` with the cyclic garbage collector held off for the duration of the capture.

    Destroying a CUDA graph (cudaGraphExecDestroy / cudaGraphDestroy) is not permitted while a stream of the process is
    capturing in the default "global" error mode.  A dead reference cycle that still owns a ``torch.cuda.CUDAGraph`` (an
    earlier trainer or model with captured steps) is finalised whenever the cyclic collector happens to run; when that is
    in the middle of a capture the capture is invalidated and the next launch reports "operation failed due to a previous
    error during capture".  So: collect BEFORE the capture begins (torch >= 2.9 no longer does), keep the collector off
    while capturing, restore it afterwards."""

    def __init__(self, graph, **kw):
        self.ctx = torch.cuda.graph(graph, **kw)

    def __enter__(self):
        import gc
        gc.collect()
        self.was_enabled = gc.isenabled()
        gc.disable()
        try:
            return self.ctx.__enter__()
        except BaseException:
            if self.was_enabled:
                gc.enable()
            raise

    def __exit__(self, *exc):
        import gc
        try:
            return self.ctx.__exit__(*exc)
        finally:
            if self.was_enabled:
                gc.enable()


_WS_SLOT = 0


class workspace_slot:
    """Convolutions enqueued inside ``with workspace_slot(k)`` use their own split-K scratch: branches of one call that
    run CONCURRENTLY on different streams (the style encoder beside the VAE encoder, the z_t decode beside the sampling
    loop) must not share the scratch that ordered launches of one stream share."""

    def __init__(self, slot: int):
        self.slot = int(slot)

    def __enter__(self):
        global _WS_SLOT
        self.old, _WS_SLOT = _WS_SLOT, self.slot
        return self

    def __exit__(self, *exc):
        global _WS_SLOT
        _WS_SLOT = self.old
        return False


def workspace(device) -> torch.Tensor:
    """Zero-initialised split-K scratch shared by all conv calls of a device and workspace slot (kernels
    leave it zeroed; launches on one stream are ordered)."""
    key = (torch.device(device).index or 0, _WS_SLOT)
    ws = _workspaces.get(key)
    if ws is None:
        ws = torch.zeros(WORKSPACE_BYTES, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


@dataclass
class Ddim:
    """Arguments of the fused DDIM update (include/mstk.h: mstk_ddim_args)."""
    x_cur: torch.Tensor
    x_next: torch.Tensor
    sab_t: float
    s1m_t: float
    sab_n: float
    s1m_n: float
    eta: float
    x_next_lp: Optional[torch.Tensor] = None
    x0_pred: Optional[torch.Tensor] = None

    def to_c(self) -> _lib.DdimArgs:
        return _lib.DdimArgs(_p(self.x_cur), _p(self.x_next), _p(self.x_next_lp), _p(self.x0_pred),
                             self.sab_t, self.s1m_t, self.sab_n, self.s1m_n, self.eta)


def conv_out_hw(h, w, ksize, stride, transposed):
    if transposed:
        return 2 * h, 2 * w
    pad = 1 if ksize == 4 else ksize // 2
    return (h + 2 * pad - ksize) // stride + 1, (w + 2 * pad - ksize) // stride + 1


# fp32 mode: contractions whose shapes fit the tcgen05 kernel run there on split (hi + lo) bf16 operands
# (csrc/split_f32.cu); False sends every fp32 contraction to the CUDA-core kernel
FP32_TC = True


def split3(src: torch.Tensor, weights: bool = False) -> torch.Tensor:
    """[..., c] fp32 -> [..., 3c] bf16: [hi | lo | hi] (activations) or [hi | hi | lo] (packed weights)"""
    lib = _lib.load()
    _dev(src, "split3.src")
    c = src.shape[-1]
    dst = torch.empty(tuple(src.shape[:-1]) + (3 * c,), dtype=torch.bfloat16, device=src.device)
    check(lib.mstk_split3(_p(src), _p(dst), src.numel() // c, c, 1 if weights else 0, _stream()), "mstk_split3")
    _count(1)
    return dst


def _conv_fp32_tc(x, wpacked, bias, *, ksize, stride, transposed, act, add, add_nc, stats, ddim, out_dtype, out_scale, out_shift, out, want_y):
    """fp32 conv on the tensor cores: split operands, one bf16 tcgen05 launch with fp32 output, residual added after"""
    w3 = getattr(wpacked, "_mstk_w3", None)
    if w3 is None:
        with torch.no_grad():
            w3 = split3(wpacked, weights=True)
        wpacked._mstk_w3 = w3                    # lives and dies with this packed copy of the parameter
    x3 = split3(x)
    y = conv(x3, w3, bias, ksize=ksize, stride=stride, transposed=transposed, act=act, add_nc=add_nc, stats=stats, ddim=ddim,
             out_dtype=out_dtype or torch.float32, out_scale=out_scale, out_shift=out_shift, out=out, want_y=want_y, backend=BACKEND_TCGEN05)
    if add is not None:
        lib = _lib.load()
        n, c = y.shape[0], y.shape[-1]
        check(lib.mstk_add_bcast(_p(y), _p(add), None, _p(y), n, y.numel() // (n * c), c, dtype_code(y), _stream()), "mstk_add_bcast")
        _count(1)
    return y


def conv(x: torch.Tensor, wpacked: torch.Tensor, bias: Optional[torch.Tensor], *, ksize: int, stride: int = 1,
         transposed: bool = False, act: int = ACT_NONE, add: Optional[torch.Tensor] = None,
         add_nc: Optional[torch.Tensor] = None, stats: Optional[torch.Tensor] = None, ddim: Optional[Ddim] = None,
         out_dtype: Optional[torch.dtype] = None, out_scale: float = 1.0, out_shift: float = 0.0,
         out: Optional[torch.Tensor] = None, want_y: bool = True, backend: int = BACKEND_AUTO) -> Optional[torch.Tensor]:
    """Conv2d / ConvTranspose2d / Linear with the fused epilogue (mstk_conv2d_fwd)."""
    lib = _lib.load()
    _dev(x, "conv.x"); _dev(wpacked, "conv.w")
    n, h, w, cin = x.shape
    kk, cout, cin_w = wpacked.shape
    if cin_w != cin or kk != ksize * ksize:
        raise MstkError(f"conv: packed weight {tuple(wpacked.shape)} does not match cin={cin} k={ksize}")
    if wpacked.dtype != x.dtype:
        raise MstkError(f"conv: weight dtype {wpacked.dtype} != activation dtype {x.dtype}")
    if (FP32_TC and backend == BACKEND_AUTO and x.dtype == torch.float32 and cin % 32 == 0 and cout % 16 == 0
            and (add is None or (want_y and add.dtype == torch.float32 and (out_dtype or x.dtype) == torch.float32))
            and tc_eligible(torch.bfloat16, 3 * cin, cout, stride, transposed, h, w)):
        return _conv_fp32_tc(x, wpacked, bias, ksize=ksize, stride=stride, transposed=transposed, act=act, add=add, add_nc=add_nc,
                             stats=stats, ddim=ddim, out_dtype=out_dtype, out_scale=out_scale, out_shift=out_shift, out=out, want_y=want_y)
    ho, wo = conv_out_hw(h, w, ksize, stride, transposed)
    out_dtype = out_dtype or x.dtype
    y = None
    if want_y:
        y = out if out is not None else torch.empty((n, ho, wo, cout), dtype=out_dtype, device=x.device)
        if tuple(y.shape) != (n, ho, wo, cout) or y.dtype != out_dtype:
            raise MstkError("conv: bad `out` tensor")
        _dev(y, "conv.out")
    if add is not None:
        _dev(add, "conv.add")
        if tuple(add.shape) != (n, ho, wo, cout) or add.dtype != x.dtype:
            raise MstkError(f"conv: add tensor {tuple(add.shape)}/{add.dtype} must match output shape and input dtype")
    if add_nc is not None and (tuple(add_nc.shape) != (n, cout) or add_nc.dtype != torch.float32):
        raise MstkError("conv: add_nc must be fp32 [n, cout]")
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != cout):
        raise MstkError("conv: bias must be fp32 [cout]")
    if stats is not None and (stats.dtype != torch.float64 or stats.numel() != 2 * cout):
        raise MstkError("conv: stats must be float64 [2*cout]")
    ws = workspace(x.device)
    dd = ddim.to_c() if ddim is not None else None
    a = _lib.ConvArgs(_p(x), _p(wpacked), _p(bias), _p(y), _p(add), _p(add_nc), _p(stats),
                      C.pointer(dd) if dd is not None else None, _p(ws), ws.numel(),
                      n, h, w, cin, cout, ksize, stride, 1 if transposed else 0, act, out_scale, out_shift,
                      dtype_code(x), dtype_code(out_dtype), backend)
    check(lib.mstk_conv2d_fwd(C.byref(a), _stream()), "mstk_conv2d_fwd")
    _count()
    if RECORD is not None:
        use_tc = backend == BACKEND_TCGEN05 or (backend == BACKEND_AUTO and tc_eligible(x.dtype, cin, cout, stride, transposed, h, w))
        keep = (x, wpacked, bias, y, add, add_nc, stats, ddim, dd, ws)

        def replay(a=a, keep=keep):
            check(lib.mstk_conv2d_fwd(C.byref(a), _stream()), "mstk_conv2d_fwd(replay)")

        RECORD.append(dict(n=n, h=h, w=w, cin=cin, cout=cout, ksize=ksize, stride=stride, transposed=transposed, note=RECORD_NOTE,
                           kernel="conv_tc_kernel" if use_tc else "conv_simt_kernel", replay=replay,
                           tag=f"{'convT' if transposed else 'conv'}k{ksize}s{stride} {cin}->{cout} @{h}x{w} n{n}"))
    return y


class ZeroPool:
    """Step-scoped pool of zeroed fp64 scratch (BatchNorm sums): ONE memset per training step instead of one fill launch per
    BatchNorm layer and direction.  The trainers activate it around their step (ZERO_POOL); slices keep their addresses
    across steps, so the captured step replays."""

    def __init__(self, device, doubles: int = 8192):
        self.buf = torch.zeros(doubles, dtype=torch.float64, device=device)
        self.off = 0

    def begin(self) -> None:
        self.buf.zero_()
        self.off = 0

    def take(self, n: int):
        span = (n + 31) // 32 * 32
        if self.off + span > self.buf.numel():
            return None
        v = self.buf[self.off:self.off + n]
        self.off += span
        return v


ZERO_POOL: Optional[ZeroPool] = None


def zeros_f64(n: int, device) -> torch.Tensor:
    """n zeroed doubles: a slice of the active step pool, else a fresh tensor"""
    if ZERO_POOL is not None and ZERO_POOL.buf.device == torch.device(device):
        v = ZERO_POOL.take(n)
        if v is not None:
            return v
    return torch.zeros(n, dtype=torch.float64, device=device)


# When a list, every un-scaled pack_weight / pack_weight_dgrad call appends what it did (source view, destination, geometry):
# the trainers turn the record of their first step into a PackPlan and refresh all packed copies with one launch per step.
PACK_RECORDER: Optional[list] = None


class PackPlan:
    """One-launch refresh of a set of packed weight copies (mstk_pack_multi).  Holds the source views and the destination
    tensors alive; the destinations keep their addresses, so module caches that hand them out stay valid."""

    def __init__(self, entries):
        lib = _lib.load()
        self.entries = list(entries)
        arr = (_lib.PackEntry * len(self.entries))()
        blocks = 0
        for i, e in enumerate(self.entries):
            arr[i] = _lib.PackEntry(_p(e["src"]), _p(e["dst"]), e["cout"], e["cin"], e["kk"], e["mode"], e["dtype"], blocks)
            blocks += lib.mstk_pack_blocks(e["cout"], e["cin"], e["kk"])
        self.blocks = blocks
        dev = self.entries[0]["dst"].device
        raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).clone()
        self.table = raw.to(dev)
        self.dst_ids = {id(e["dst"]) for e in self.entries}

    def __len__(self):
        return len(self.entries)

    def run(self) -> None:
        lib = _lib.load()
        check(lib.mstk_pack_multi(_p(self.table), len(self.entries), self.blocks, _stream()), "mstk_pack_multi")
        _count(1)


def pack_weight(w: torch.Tensor, dtype: torch.dtype, *, transposed: bool = False,
                scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    """nn.Conv2d [co,ci,k,k] / nn.ConvTranspose2d [ci,co,k,k] / nn.Linear [co,ci] fp32 -> [k*k, co, ci]."""
    lib = _lib.load()
    w = _dev(w.detach(), "pack.w")
    if w.dtype != torch.float32:
        raise MstkError("pack_weight: parameters must be fp32")
    if w.dim() == 2:
        co, ci, k = w.shape[0], w.shape[1], 1
    elif transposed:
        ci, co, k = w.shape[0], w.shape[1], w.shape[2]
    else:
        co, ci, k = w.shape[0], w.shape[1], w.shape[2]
    dst = torch.empty((k * k, co, ci), dtype=dtype, device=w.device)
    if PACK_RECORDER is not None and scale is None:
        PACK_RECORDER.append(dict(src=w, dst=dst, cout=co, cin=ci, kk=k * k, mode=1 if (transposed and w.dim() == 4) else 0,
                                  dtype=dtype_code(dtype)))
    fn = lib.mstk_pack_convt_weight if (transposed and w.dim() == 4) else lib.mstk_pack_conv_weight
    if transposed and w.dim() == 4:
        check(fn(_p(w), _p(scale), _p(dst), ci, co, k, dtype_code(dtype), _stream()), "mstk_pack_convt_weight")
    else:
        check(fn(_p(w), _p(scale), _p(dst), co, ci, k, dtype_code(dtype), _stream()), "mstk_pack_conv_weight")
    return dst


def nchw_to_nhwc(x: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    lib = _lib.load()
    x = _dev(x, "nchw_to_nhwc.x")
    if x.dtype != torch.float32:
        raise MstkError("nchw_to_nhwc: source must be fp32")
    n, c, h, w = x.shape
    dst = torch.empty((n, h, w, c), dtype=dtype, device=x.device)
    check(lib.mstk_nchw_to_nhwc(_p(x), _p(dst), n, c, h, w, dtype_code(dtype), _stream()), "mstk_nchw_to_nhwc")
    _count(1)
    return dst


def nhwc_to_nchw(x: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    x = _dev(x, "nhwc_to_nchw.x")
    n, h, w, c = x.shape
    dst = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device)
    check(lib.mstk_nhwc_to_nchw(_p(x), _p(dst), n, c, h, w, dtype_code(x), _stream()), "mstk_nhwc_to_nchw")
    _count(1)
    return dst


def u8_gather_crop(pool: torch.Tensor, index: torch.Tensor, crop_h: int, crop_w: int) -> torch.Tensor:
    """pool [n,h,w] uint8 (CUDA), index [b] int64 (CUDA) -> [b,1,crop_h,crop_w] fp32 in [0,1] (top-left crop / 255)"""
    lib = _lib.load()
    _dev(pool, "u8_gather_crop.pool"); _dev(index, "u8_gather_crop.index")
    if pool.dtype != torch.uint8 or pool.dim() != 3 or index.dtype != torch.int64:
        raise MstkError("u8_gather_crop: pool must be uint8 [n,h,w] and index int64 [b]")
    b = index.numel()
    out = torch.empty((b, 1, crop_h, crop_w), dtype=torch.float32, device=pool.device)
    check(lib.mstk_u8_gather_crop(_p(pool), _p(index), _p(out), b, pool.shape[1], pool.shape[2], crop_h, crop_w, _stream()), "mstk_u8_gather_crop")
    _count(1)
    return out


def cast(x: torch.Tensor, dtype: torch.dtype, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    lib = _lib.load()
    x = _dev(x, "cast.x")
    if x.dtype == dtype and out is None:
        return x
    dst = out if out is not None else torch.empty(x.shape, dtype=dtype, device=x.device)
    if dst.dtype != dtype or dst.numel() != x.numel() or not dst.is_contiguous():
        raise MstkError("cast: bad `out` tensor")
    check(lib.mstk_cast(_p(x), dtype_code(x), _p(dst), dtype_code(dtype), x.numel(), _stream()), "mstk_cast")
    _count(1)
    return dst


def bn_apply(x: torch.Tensor, stats: torch.Tensor, gamma, beta, *, count: int, eps: float, momentum: float,
             act: int, out_dtype: torch.dtype, running_mean=None, running_var=None, save_mean=None,
             save_invstd=None) -> torch.Tensor:
    lib = _lib.load()
    x = _dev(x, "bn.x")
    c = x.shape[-1]
    y = torch.empty(x.shape, dtype=out_dtype, device=x.device)
    a = _lib.BnArgs(_p(x), _p(y), _p(stats), _p(gamma), _p(beta), _p(running_mean), _p(running_var), _p(save_mean),
                    _p(save_invstd), count, x.numel() // c, c, eps, momentum, act, dtype_code(x), dtype_code(out_dtype))
    check(lib.mstk_bn_apply(C.byref(a), _stream()), "mstk_bn_apply")
    _count(1)
    return y


def time_embed(t: torch.Tensor, w1, b1, w2, b2) -> torch.Tensor:
    lib = _lib.load()
    t = _dev(t, "time_embed.t")
    if t.dtype != torch.float32:
        raise MstkError("time_embed: t must be fp32 (cast on the host side)")
    dim = w1.shape[0]
    out = torch.empty((t.numel(), dim), dtype=torch.float32, device=t.device)
    check(lib.mstk_time_embed(_p(t), t.numel(), _p(w1), _p(b1), _p(w2), _p(b2), _p(out), dim, _stream()),
          "mstk_time_embed")
    _count(1)
    return out


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int) -> torch.Tensor:
    """q [b,l,e]; k, v [b,s,e] (last-dim-contiguous views allowed) -> [b,l,e]."""
    lib = _lib.load()
    b, l, e = q.shape
    s = k.shape[1]
    for t, nm in ((q, "q"), (k, "k"), (v, "v")):
        if not t.is_cuda or t.stride(-1) != 1 or t.stride(0) != t.shape[1] * t.stride(1):
            raise MstkError(f"attention: {nm} must be a CUDA tensor with dense tokens")
    out = torch.empty((b, l, e), dtype=q.dtype, device=q.device)
    a = _lib.AttnArgs(_p(q), _p(k), _p(v), _p(out), b, l, s, e, heads, q.stride(1), k.stride(1), v.stride(1), e,
                      dtype_code(q))
    check(lib.mstk_attention_fwd(C.byref(a), _stream()), "mstk_attention_fwd")
    _count(1)
    if RECORD is not None:
        keep = (q, k, v, out)

        def replay(a=a, keep=keep):
            check(lib.mstk_attention_fwd(C.byref(a), _stream()), "mstk_attention_fwd(replay)")

        RECORD.append(dict(kernel="attention_block_kernel", replay=replay, flops=4.0 * b * l * s * e, note=RECORD_NOTE,
                           tag=f"attention core b{b} l{l} s{s} e{e} h{heads}"))
    return out


def xattn_compose(in_proj_weight: torch.Tensor, in_proj_bias: torch.Tensor, out_proj_weight: torch.Tensor, heads: int,
                  dtype: torch.dtype):
    """Composed projection of the style side of a cross-attention layer (csrc/xattn.cu): packed k1 conv weight
    [1, ctot, e] in ``dtype`` and fp32 bias [ctot]."""
    lib = _lib.load()
    w_in, b_in, w_out = (_dev(t.detach(), "xattn_compose") for t in (in_proj_weight, in_proj_bias, out_proj_weight))
    if any(t.dtype != torch.float32 for t in (w_in, b_in, w_out)):
        raise MstkError("xattn_compose: parameters must be fp32")
    e = w_out.shape[0]
    ctot = lib.mstk_xattn_ctot(e, heads)
    wp = torch.empty((1, ctot, e), dtype=dtype, device=w_in.device)
    bias = torch.empty(ctot, dtype=torch.float32, device=w_in.device)
    check(lib.mstk_xattn_compose(_p(w_in), _p(b_in), _p(w_out), _p(wp), _p(bias), e, heads, dtype_code(dtype), _stream()), "mstk_xattn_compose")
    return wp, bias


def xattn(x: torch.Tensor, kvp: torch.Tensor, bias: Optional[torch.Tensor], heads: int, kv_static: bool = False) -> torch.Tensor:
    """x [b,l,e], kvp [b,s,ctot] (style tokens through the composed projection) -> [b,l,e]: the whole cross-attention layer."""
    lib = _lib.load()
    _dev(x, "xattn.x"); _dev(kvp, "xattn.kvp")
    b, l, e = x.shape
    s, ctot = kvp.shape[1], kvp.shape[2]
    if kvp.shape[0] != b or kvp.dtype != x.dtype:
        raise MstkError("xattn: kvp must match x in batch and dtype")
    out = torch.empty((b, l, e), dtype=x.dtype, device=x.device)
    a = _lib.XattnArgs(_p(x), _p(kvp), _p(bias), _p(out), b, l, s, e, heads, ctot, 1 if kv_static else 0, dtype_code(x))
    check(lib.mstk_xattn_fwd(C.byref(a), _stream()), "mstk_xattn_fwd")
    _count(1)
    if RECORD is not None:
        keep = (x, kvp, bias, out)

        def replay(a=a, keep=keep):
            check(lib.mstk_xattn_fwd(C.byref(a), _stream()), "mstk_xattn_fwd(replay)")

        # algorithmic work of the layer as the reference states it: q and out projections + the attention core
        RECORD.append(dict(kernel="xattn_kernel", replay=replay, flops=4.0 * b * l * e * e + 4.0 * b * l * s * e, note=RECORD_NOTE,
                           tag=f"cross-attention (composed) b{b} l{l} s{s} e{e} h{heads}"))
    return out


def q_sample(x0, eps, alpha_bar, t, *, lp: bool = False):
    lib = _lib.load()
    _dev(x0, "q_sample.x0"); _dev(eps, "q_sample.eps")
    b = x0.shape[0]
    inner = x0.numel() // b
    z = torch.empty_like(x0)
    zlp = torch.empty(x0.shape, dtype=torch.bfloat16, device=x0.device) if lp else None
    check(lib.mstk_q_sample(_p(x0), _p(eps), _p(alpha_bar), _p(t), _p(z), _p(zlp), b, inner, _stream()), "mstk_q_sample")
    _count(1)
    return z, zlp


def predict_x0(z_t, eps_hat, alpha_bar, t, *, lp: bool = False):
    lib = _lib.load()
    _dev(z_t, "predict_x0.z_t"); _dev(eps_hat, "predict_x0.eps_hat")
    b = z_t.shape[0]
    inner = z_t.numel() // b
    x0 = torch.empty_like(z_t)
    xlp = torch.empty(z_t.shape, dtype=torch.bfloat16, device=z_t.device) if lp else None
    check(lib.mstk_predict_x0(_p(z_t), _p(eps_hat), _p(alpha_bar), _p(t), _p(x0), _p(xlp), b, inner, _stream()),
          "mstk_predict_x0")
    _count(1)
    return x0, xlp


def ddim_step(d: Ddim, eps: torch.Tensor) -> None:
    lib = _lib.load()
    dd = d.to_c()
    check(lib.mstk_ddim_step(C.byref(dd), _p(_dev(eps, "ddim.eps")), eps.numel(), _stream()), "mstk_ddim_step")
    _count(1)


_scratch = {}


def _loss_scratch(device, n_doubles: int) -> torch.Tensor:
    key = (torch.device(device).index or 0, n_doubles, _WS_SLOT)        # concurrent branches (workspace_slot) get their own
    s = _scratch.get(key)
    if s is None:
        s = torch.zeros(n_doubles, dtype=torch.float64, device=device)
        _scratch[key] = s
    return s


def mse(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    _dev(a, "mse.a"); _dev(b, "mse.b")
    if a.dtype != b.dtype or a.numel() != b.numel():
        raise MstkError("mse: operands must match")
    out = torch.empty(1, dtype=torch.float32, device=a.device)
    check(lib.mstk_mse(_p(a), _p(b), dtype_code(a), a.numel(), _p(_loss_scratch(a.device, 8)), _p(out), _stream()), "mstk_mse")
    _count(1)
    return out


def kl(z: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    _dev(z, "kl.z")
    out = torch.empty(1, dtype=torch.float32, device=z.device)
    check(lib.mstk_kl(_p(z), z.numel(), _p(_loss_scratch(z.device, 8)), _p(out), _stream()), "mstk_kl")
    _count(1)
    return out


def feature_std_mse(f: torch.Tensor, g: torch.Tensor, weight: float, out: torch.Tensor) -> None:
    """Accumulates weight * mean((f/(std_i(f)+1e-8) - g/(std_i(g)+1e-8))^2) into out[0]."""
    lib = _lib.load()
    _dev(f, "feat.f"); _dev(g, "feat.g")
    b = f.shape[0]
    inner = f.numel() // b
    tmp = _loss_scratch(f.device, 4 * b + 4)
    check(lib.mstk_feature_std_mse(_p(f), _p(g), dtype_code(f), b, inner, _p(tmp), weight, _p(out), _stream()),
          "mstk_feature_std_mse")
    _count(2)


def feature_mse3(f_all: torch.Tensor, b: int, weight: float, out_a: torch.Tensor, out_c: torch.Tensor) -> None:
    """f_all: [3b, ...] features of three image batches stacked along dim 0.  Accumulates the VGGish tap term
    (loss.py:94-99) of the pair (f_all[:b], f_all[b:2b]) into out_a and of (f_all[b:2b], f_all[2b:]) into out_c."""
    lib = _lib.load()
    _dev(f_all, "feat3.f")
    if f_all.shape[0] != 3 * b or not f_all.is_contiguous():
        raise MstkError("feature_mse3: expected a contiguous [3b, ...] tensor")
    inner = f_all.numel() // (3 * b)
    esz = f_all.element_size()
    tmp = _loss_scratch(f_all.device, 6 * b + 4)
    dt = dtype_code(f_all)
    base = f_all.data_ptr()
    check(lib.mstk_sample_stats(base, dt, 3 * b, inner, _p(tmp), _stream()), "mstk_sample_stats")
    d0 = tmp.data_ptr()
    check(lib.mstk_feature_mse3(base, base + b * inner * esz, base + 2 * b * inner * esz, dt, b, inner, d0, d0 + 16 * b, d0 + 32 * b,
                                d0 + 48 * b, weight, _p(out_a), _p(out_c), _stream()), "mstk_feature_mse3")
    _count(2)


def feature_fused3(f_all: torch.Tensor, b: int, weight: float, out_a: torch.Tensor, out_c: torch.Tensor, pool: bool):
    """One pass over a VGGish tap f_all [3b,h,w,c]: accumulates the tap terms of (f[:b], f[b:2b]) into out_a and of
    (f[b:2b], f[2b:]) into out_c; returns the 2x2 max-pooled map when ``pool`` (else None)."""
    lib = _lib.load()
    _dev(f_all, "feat_fused3.f")
    n3, h, w, c = f_all.shape
    if n3 != 3 * b:
        raise MstkError("feature_fused3: expected a [3b,h,w,c] tensor")
    pooled = torch.empty((n3, h // 2, w // 2, c), dtype=f_all.dtype, device=f_all.device) if pool else None
    tmp = _loss_scratch(f_all.device, 8 * b + 2)
    check(lib.mstk_feature_fused3(_p(f_all), _p(pooled), dtype_code(f_all), b, h, w, c, tmp.data_ptr(), tmp.data_ptr() + 64 * b, weight,
                                  _p(out_a), _p(out_c), _stream()), "mstk_feature_fused3")
    _count(1)
    return pooled


def feature_first3(three: torch.Tensor, wpacked: torch.Tensor, bias: torch.Tensor, b: int, weight: float, out_a: torch.Tensor,
                   out_c: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """First VGGish tap of the training-step loss without materialising it: three [3b,h,w,1] fp32 images (x; recon; style),
    features.0's packed fp32 weight [9,64,1] and bias -> accumulates the tap terms into out_a / out_c and returns the 2x2
    max-pooled map [3b,h/2,w/2,64] in ``dtype`` (the next conv's input)."""
    lib = _lib.load()
    _dev(three, "feature_first3.x"); _dev(wpacked, "feature_first3.w")
    n3, h, w, c1 = three.shape
    if n3 != 3 * b or c1 != 1 or three.dtype != torch.float32 or tuple(wpacked.shape) != (9, 64, 1) or wpacked.dtype != torch.float32:
        raise MstkError("feature_first3: expected [3b,h,w,1] fp32 images and a packed fp32 [9,64,1] weight")
    pooled = torch.empty((n3, h // 2, w // 2, 64), dtype=dtype, device=three.device)
    tmp = _loss_scratch(three.device, 8 * b + 2)
    check(lib.mstk_feature_first3(_p(three), _p(wpacked), _p(bias), _p(pooled), dtype_code(dtype), b, h, w, tmp.data_ptr(), tmp.data_ptr() + 64 * b,
                                  weight, _p(out_a), _p(out_c), _stream()), "mstk_feature_first3")
    _count(1)
    return pooled


def maxpool2x2(x: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    _dev(x, "maxpool.x")
    n, h, w, c = x.shape
    y = torch.empty((n, h // 2, w // 2, c), dtype=x.dtype, device=x.device)
    check(lib.mstk_maxpool2x2(_p(x), _p(y), n, h, w, c, dtype_code(x), _stream()), "mstk_maxpool2x2")
    _count(1)
    return y


def mel_band_ranges(mel_fb: torch.Tensor) -> torch.Tensor:
    """[n_mels, 2] int32: first / one-past-last non-zero bin of every band of a filterbank (once per filterbank)"""
    lib = _lib.load()
    _dev(mel_fb, "mel.fb")
    r = torch.empty((mel_fb.shape[0], 2), dtype=torch.int32, device=mel_fb.device)
    check(lib.mstk_mel_band_ranges(_p(mel_fb), mel_fb.shape[0], _p(r), _stream()), "mstk_mel_band_ranges")
    return r


def mel_frontend(wave: torch.Tensor, mel_fb: torch.Tensor, crop_frames: int, want_db: bool = False, band_ranges: Optional[torch.Tensor] = None):
    lib = _lib.load()
    _dev(wave, "mel.wave"); _dev(mel_fb, "mel.fb")
    b, ln = wave.shape
    n_mels = mel_fb.shape[0]
    frames = 1 + ln // 512
    out = torch.empty((b, n_mels, crop_frames), dtype=torch.float32, device=wave.device)
    db = torch.empty((b, n_mels, frames), dtype=torch.float32, device=wave.device) if want_db else None
    nbytes = lib.mstk_mel_workspace_bytes(b, ln, n_mels)
    work = torch.empty(nbytes, dtype=torch.uint8, device=wave.device)
    a = _lib.MelArgs(_p(wave), _p(out), _p(db), _p(mel_fb), _p(work), nbytes, b, ln, n_mels, crop_frames, _p(band_ranges))
    check(lib.mstk_mel_frontend(C.byref(a), _stream()), "mstk_mel_frontend")
    _count(2 if band_ranges is not None else 3)
    return out, db


def mel_to_stft(image_u8: torch.Tensor, pinv: torch.Tensor, mel_fb: torch.Tensor, band_ranges: torch.Tensor, *, max_db: float = 80.0,
                pg_iters: int = 0, pg_step: float = 0.0):
    """image [b, n_mels, frames] uint8 -> (mag [b, frames, 1025] fp32, pg_norm [b] fp32): librosa.feature.inverse.mel_to_stft at
    the NNLS starting point (audio_processor.py:93-97 + mel_to_audio's first half)."""
    lib = _lib.load()
    _dev(image_u8, "mel_to_stft.image"); _dev(pinv, "mel_to_stft.pinv"); _dev(mel_fb, "mel_to_stft.fb"); _dev(band_ranges, "mel_to_stft.ranges")
    if image_u8.dtype != torch.uint8 or image_u8.dim() != 3:
        raise MstkError(f"mel_to_stft: image must be uint8 [b, n_mels, frames], got {image_u8.dtype} {tuple(image_u8.shape)}")
    b, n_mels, frames = image_u8.shape
    if tuple(pinv.shape) != (1025, n_mels) or tuple(mel_fb.shape) != (n_mels, 1025):
        raise MstkError("mel_to_stft: pinv must be [1025, n_mels] and mel_fb [n_mels, 1025]")
    mag = torch.empty((b, frames, 1025), dtype=torch.float32, device=image_u8.device)
    pg = torch.empty((b,), dtype=torch.float32, device=image_u8.device)
    a = _lib.MelToStftArgs(_p(image_u8), _p(pinv), _p(mel_fb), _p(band_ranges), _p(mag), _p(pg), b, n_mels, frames, float(max_db),
                           int(pg_iters), float(pg_step))
    check(lib.mstk_mel_to_stft(C.byref(a), _stream()), "mstk_mel_to_stft")
    _count(1)
    return mag, pg


def db_image_to_mag(image_u8: torch.Tensor, *, max_db: float = 80.0) -> torch.Tensor:
    """image [b, bins, frames] uint8 dB -> amplitude [b, frames, bins] fp32 (audio_processor.py:150-155)."""
    lib = _lib.load()
    _dev(image_u8, "db_image_to_mag.image")
    if image_u8.dtype != torch.uint8 or image_u8.dim() != 3:
        raise MstkError(f"db_image_to_mag: image must be uint8 [b, bins, frames], got {image_u8.dtype} {tuple(image_u8.shape)}")
    b, bins, frames = image_u8.shape
    mag = torch.empty((b, frames, bins), dtype=torch.float32, device=image_u8.device)
    check(lib.mstk_db_image_to_mag(_p(image_u8), _p(mag), b, bins, frames, float(max_db), _stream()), "mstk_db_image_to_mag")
    _count(1)
    return mag


def griffinlim(mag: torch.Tensor, phase_u: torch.Tensor, *, n_iter: int = 32, momentum: float = 0.99) -> torch.Tensor:
    """mag, phase_u [b, frames, 1025] fp32 -> audio [b, 512 * (frames - 1)] fp32 (librosa.griffinlim, n_fft 2048, hop 512)."""
    lib = _lib.load()
    _dev(mag, "griffinlim.mag"); _dev(phase_u, "griffinlim.phase_u")
    if mag.dim() != 3 or mag.shape[2] != 1025 or mag.shape != phase_u.shape or mag.dtype != torch.float32 or phase_u.dtype != torch.float32:
        raise MstkError(f"griffinlim: mag / phase_u must be fp32 [b, frames, 1025], got {tuple(mag.shape)} / {tuple(phase_u.shape)}")
    b, frames, _ = mag.shape
    if frames < 2:
        raise MstkError("griffinlim: at least 2 frames")
    audio = torch.empty((b, 512 * (frames - 1)), dtype=torch.float32, device=mag.device)
    nbytes = lib.mstk_griffinlim_workspace_bytes(b, frames)
    work = torch.empty(nbytes, dtype=torch.uint8, device=mag.device)
    a = _lib.GriffinLimArgs(_p(mag), _p(phase_u), _p(audio), _p(work), nbytes, b, frames, int(n_iter), float(momentum))
    check(lib.mstk_griffinlim(C.byref(a), _stream()), "mstk_griffinlim")
    _count(4 + 2 * int(n_iter))
    return audio


# ------------------------------------------------------------------------------------------
# backward pass / optimizer (include/mstk.h, "Backward pass and optimizer" section)
# ------------------------------------------------------------------------------------------
def pack_weight_dgrad(w: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """Conv2d [co,ci,k,k] / Linear [co,ci] fp32 -> [k*k, ci, co] with flipped taps: the packed weight of the
    stride-1 conv that maps dL/dy to dL/dx."""
    lib = _lib.load()
    w = _dev(w.detach(), "pack_dgrad.w")
    co, ci = w.shape[0], w.shape[1]
    k = 1 if w.dim() == 2 else w.shape[2]
    dst = torch.empty((k * k, ci, co), dtype=dtype, device=w.device)
    if PACK_RECORDER is not None:
        PACK_RECORDER.append(dict(src=w, dst=dst, cout=co, cin=ci, kk=k * k, mode=2, dtype=dtype_code(dtype)))
    check(lib.mstk_pack_conv_weight_dgrad(_p(w), _p(dst), co, ci, k, dtype_code(dtype), _stream()), "mstk_pack_conv_weight_dgrad")
    return dst


def wgrad(x: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor, *, ksize: int, stride: int = 1, transposed: bool = False) -> None:
    """dw (fp32, the parameter's own layout) += weight gradient of the forward conv that mapped x -> y."""
    lib = _lib.load()
    _dev(x, "wgrad.x"); _dev(dy, "wgrad.dy"); _dev(dw, "wgrad.dw")
    n, h, w, cin = x.shape
    cout = dy.shape[-1]
    ho, wo = conv_out_hw(h, w, ksize, stride, transposed)
    if tuple(dy.shape) != (n, ho, wo, cout):
        raise MstkError(f"wgrad: dy {tuple(dy.shape)} does not match the forward output {(n, ho, wo, cout)}")
    if dw.dtype != torch.float32 or dw.numel() != cin * cout * ksize * ksize:
        raise MstkError("wgrad: dw must be the fp32 parameter-shaped gradient")
    a = _lib.WgradArgs(_p(x), _p(dy), _p(dw), n, h, w, cin, cout, ksize, stride, 1 if transposed else 0, dtype_code(x), dtype_code(dy))
    check(lib.mstk_conv2d_wgrad(C.byref(a), _stream()), "mstk_conv2d_wgrad")
    _count()


def act_bwd(dy: torch.Tensor, y: Optional[torch.Tensor], *, act: int, dy2: Optional[torch.Tensor] = None,
            dbias: Optional[torch.Tensor] = None, g_dtype: Optional[torch.dtype] = None, want_g: bool = True,
            y_scale: float = 1.0, y_shift: float = 0.0) -> Optional[torch.Tensor]:
    lib = _lib.load()
    _dev(dy, "act_bwd.dy")
    c = dy.shape[-1]
    rows = dy.numel() // c
    g = torch.empty(dy.shape, dtype=g_dtype or dy.dtype, device=dy.device) if want_g else None
    if dy2 is not None and (dy2.dtype != dy.dtype or dy2.shape != dy.shape):
        raise MstkError("act_bwd: dy2 must match dy")
    if y is not None and y.numel() != dy.numel():
        raise MstkError("act_bwd: y must match dy")
    a = _lib.ActBwdArgs(_p(dy), _p(dy2), _p(y), _p(g), _p(dbias), rows, c, act, y_scale, y_shift, dtype_code(dy),
                        dtype_code(y) if y is not None else 0, dtype_code(g) if g is not None else 0)
    check(lib.mstk_act_bwd(C.byref(a), _stream()), "mstk_act_bwd")
    _count()
    return g


def bn_bwd(dy: torch.Tensor, v: torch.Tensor, mean, invstd, gamma, beta, *, act: int, dgamma, dbeta, world: int = 1,
           sync=None, dv_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """training-mode BatchNorm2d backward (two passes; `sync(sums)` all-reduces the two per-channel sums)"""
    lib = _lib.load()
    c = v.shape[-1]
    rows = v.numel() // c
    sums = zeros_f64(2 * c, v.device)
    dv = torch.empty(v.shape, dtype=dv_dtype or dy.dtype, device=v.device)
    a = _lib.BnBwdArgs(_p(dy), _p(v), _p(dv), _p(mean), _p(invstd), _p(gamma), _p(beta), _p(sums), _p(dgamma), _p(dbeta),
                       rows, rows * world, c, act, dtype_code(dy), dtype_code(v), dtype_code(dv))
    check(lib.mstk_bn_bwd_reduce(C.byref(a), _stream()), "mstk_bn_bwd_reduce")
    if sync is not None:
        sync(sums)
    check(lib.mstk_bn_bwd_apply(C.byref(a), _stream()), "mstk_bn_bwd_apply")
    _count(2)
    return dv


def attention_bwd(q, k, v, d_out, heads: int):
    """q, d_out [b,l,e]; k, v [b,s,e] views -> dq [b,l,e], dkv [b,s,2e] (dk = [..., :e], dv = [..., e:])"""
    lib = _lib.load()
    b, l, e = q.shape
    s = k.shape[1]
    dq = torch.empty((b, l, e), dtype=q.dtype, device=q.device)
    dkv = torch.empty((b, s, 2 * e), dtype=q.dtype, device=q.device)
    a = _lib.AttnBwdArgs(_p(q), _p(k), _p(v), _p(d_out), _p(dq), dkv.data_ptr(), dkv.data_ptr() + e * dkv.element_size(),
                         b, l, s, e, heads, q.stride(1), k.stride(1), v.stride(1), d_out.stride(1), e, 2 * e, 2 * e, dtype_code(q))
    check(lib.mstk_attention_bwd(C.byref(a), _stream()), "mstk_attention_bwd")
    _count()
    return dq, dkv


def time_embed_train(t, w1, b1, w2, b2):
    lib = _lib.load()
    dim = w1.shape[0]
    n = t.numel()
    out, e, pre, h = (torch.empty((n, dim), dtype=torch.float32, device=t.device) for _ in range(4))
    check(lib.mstk_time_embed_train(_p(t), n, _p(w1), _p(b1), _p(w2), _p(b2), _p(out), _p(e), _p(pre), _p(h), dim, _stream()),
          "mstk_time_embed_train")
    _count()
    return out, e, pre, h


def gelu_bwd(dh, pre):
    lib = _lib.load()
    out = torch.empty_like(pre)
    check(lib.mstk_gelu_bwd(_p(dh), _p(pre), _p(out), pre.numel(), _stream()), "mstk_gelu_bwd")
    _count()
    return out


def sum_hw(a: torch.Tensor, a2: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[b,h,w,c] -> fp32 [b,c] (sum of a + a2 over the pixels)"""
    lib = _lib.load()
    b, c = a.shape[0], a.shape[-1]
    hw = a.numel() // (b * c)
    out = torch.empty((b, c), dtype=torch.float32, device=a.device)
    check(lib.mstk_sum_hw(_p(a), _p(a2), dtype_code(a), b, hw, c, _p(out), _stream()), "mstk_sum_hw")
    _count()
    return out


def add_bcast(y: torch.Tensor, add: Optional[torch.Tensor] = None, add_nc: Optional[torch.Tensor] = None) -> torch.Tensor:
    lib = _lib.load()
    n, c = y.shape[0], y.shape[-1]
    hw = y.numel() // (n * c)
    out = torch.empty_like(y)
    check(lib.mstk_add_bcast(_p(y), _p(add), _p(add_nc), _p(out), n, hw, c, dtype_code(y), _stream()), "mstk_add_bcast")
    _count()
    return out


def scale_dev_(x: torch.Tensor, factor: torch.Tensor) -> torch.Tensor:
    """x (fp32, in place) *= factor (a 1-element fp32 CUDA tensor)"""
    lib = _lib.load()
    _dev(x, "scale_dev.x")
    if x.dtype != torch.float32 or factor.dtype != torch.float32 or factor.numel() != 1 or not factor.is_cuda:
        raise MstkError("scale_dev_: fp32 tensor and a 1-element fp32 CUDA factor expected")
    check(lib.mstk_scale_dev(_p(x), _p(factor), x.numel(), _stream()), "mstk_scale_dev")
    _count()
    return x


def mse_grad(a: torch.Tensor, b: torch.Tensor, scale: float, g_dtype: torch.dtype) -> torch.Tensor:
    lib = _lib.load()
    g = torch.empty(a.shape, dtype=g_dtype, device=a.device)
    check(lib.mstk_mse_grad(_p(a), _p(b), scale, _p(g), dtype_code(g_dtype), a.numel(), _stream()), "mstk_mse_grad")
    _count()
    return g


def eps_grad(dx0: torch.Tensor, eps_hat, eps, alpha_bar, t, scale_d: float, g_dtype: torch.dtype) -> torch.Tensor:
    lib = _lib.load()
    b = eps_hat.shape[0]
    g = torch.empty(eps_hat.shape, dtype=g_dtype, device=eps_hat.device)
    check(lib.mstk_eps_grad(_p(dx0), dtype_code(dx0), _p(eps_hat), _p(eps), _p(alpha_bar), _p(t), scale_d, _p(g), dtype_code(g_dtype),
                            b, eps_hat.numel() // b, _stream()), "mstk_eps_grad")
    _count()
    return g


def kl_grad(z: torch.Tensor, dz: Optional[torch.Tensor], scale: float) -> torch.Tensor:
    """fp32 g = dz + scale * (z - z/(z^2 + 1e-8)); scale = weight / z.numel() for weight * kl_regularization_loss(z)"""
    lib = _lib.load()
    _dev(z, "kl_grad.z")
    if z.dtype != torch.float32 or (dz is not None and dz.shape != z.shape):
        raise MstkError("kl_grad: z must be fp32 and dz must match it")
    g = torch.empty_like(z)
    check(lib.mstk_kl_grad(_p(z), _p(dz), dtype_code(dz) if dz is not None else 0, scale, _p(g), z.numel(), _stream()), "mstk_kl_grad")
    _count()
    return g


def adam_step(p, g, m, v, *, lr: float, beta1: float, beta2: float, eps: float, step, grad_scale: float = 1.0,
              weight_decay: Optional[float] = None) -> None:
    """``step``: python int, or a 1-element int64 CUDA tensor holding the step count (graph-replayable).
    ``weight_decay`` not None selects torch.optim.AdamW's decoupled decay (device step counter only)."""
    lib = _lib.load()
    if weight_decay is not None:
        if not isinstance(step, torch.Tensor) or step.dtype != torch.int64 or not step.is_cuda:
            raise MstkError("adam_step: AdamW needs an int64 CUDA step counter")
        check(lib.mstk_adamw_step_dev(_p(p), _p(g), _p(m), _p(v), p.numel(), lr, beta1, beta2, eps, weight_decay, _p(step), grad_scale,
                                      _stream()), "mstk_adamw_step_dev")
        _count()
        return
    if isinstance(step, torch.Tensor):
        if step.dtype != torch.int64 or not step.is_cuda:
            raise MstkError("adam_step: a device step counter must be an int64 CUDA tensor")
        check(lib.mstk_adam_step_dev(_p(p), _p(g), _p(m), _p(v), p.numel(), lr, beta1, beta2, eps, _p(step), grad_scale, _stream()),
              "mstk_adam_step_dev")
        _count()
        return
    check(lib.mstk_adam_step(_p(p), _p(g), _p(m), _p(v), p.numel(), lr, beta1, beta2, eps, step, grad_scale, _stream()), "mstk_adam_step")
    _count()
