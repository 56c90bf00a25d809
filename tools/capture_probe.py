"""Scratch diagnostic: runs a pytest selection with a C-call profiler that reports the first C function after whose return the
current stream's capture is found invalidated (cuStreamIsCapturing status 2).  usage: python tools/capture_probe.py <pytest args>"""
import ctypes
import sys
import traceback

import pytest
import torch

cu = ctypes.CDLL("libcuda.so.1")
state = {"done": False, "last_ok": None}


def prof(frame, event, arg):
    if state["done"] or event not in ("c_return", "c_exception"):
        return
    if not torch.cuda.is_initialized():
        return
    s = torch.cuda.current_stream().cuda_stream
    if s == 0:
        return
    st = ctypes.c_int(0)
    rc = cu.cuStreamIsCapturing(ctypes.c_void_p(s), ctypes.byref(st))
    if st.value == 2 or rc == 901:
        state["done"] = True
        sys.stderr.write(f"\n#### capture INVALIDATED after C call {arg!r} (rc={rc}); previous clean C call: {state['last_ok']!r}\n")
        traceback.print_stack(frame, file=sys.stderr)
        sys.stderr.write("####\n")
    elif st.value == 1:
        state["last_ok"] = getattr(arg, "__qualname__", repr(arg))


sys.setprofile(prof)
rc = pytest.main(sys.argv[1:])
sys.setprofile(None)
sys.exit(int(rc))
