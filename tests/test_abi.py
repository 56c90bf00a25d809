"""CPU: the C-ABI library loads and exports every symbol include/mstk.h declares, and the ctypes
binding table covers exactly that set (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "mstk.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mstk_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    from music_style_transfer_ldm_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in mstk.h but not exported: {missing}"


def test_binding_table_matches_header():
    from music_style_transfer_ldm_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    lib = _lib.load()
    assert lib.mstk_version() == 1
    assert lib.mstk_last_error() is not None


def test_struct_sizes_match_the_header_layout(tmp_path):
    # sizeof() as the C compiler sees include/mstk.h vs the ctypes mirrors
    import subprocess
    from music_style_transfer_ldm_b200 import _lib
    pairs = [("mstk_ddim_args", _lib.DdimArgs), ("mstk_conv_args", _lib.ConvArgs), ("mstk_bn_args", _lib.BnArgs),
             ("mstk_attn_args", _lib.AttnArgs), ("mstk_mel_args", _lib.MelArgs), ("mstk_wgrad_args", _lib.WgradArgs),
             ("mstk_act_bwd_args", _lib.ActBwdArgs), ("mstk_bn_bwd_args", _lib.BnBwdArgs), ("mstk_attn_bwd_args", _lib.AttnBwdArgs),
             ("mstk_mel_to_stft_args", _lib.MelToStftArgs), ("mstk_griffinlim_args", _lib.GriffinLimArgs),
             ("mstk_pack_entry", _lib.PackEntry), ("mstk_peer_ctx", _lib.PeerCtx)]
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "mstk.h"\nint main(void){' +
                   "".join(f'printf("%zu\\n", sizeof({c}));' for c, _ in pairs) + "return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    for (cname, cls), sz in zip(pairs, sizes):
        assert ctypes.sizeof(cls) == sz, f"{cname}: header {sz} bytes, ctypes {ctypes.sizeof(cls)}"
