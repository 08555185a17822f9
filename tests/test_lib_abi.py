"""CPU: the C-ABI library loads and exports every symbol include/sat_b200.h declares; struct sizes agree."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "sat_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sat_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from sat_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import importlib.util
        spec = importlib.util.spec_from_file_location("sat_build", os.path.join(ROOT, "show-attend-and-tell_b200", "build.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    names = _header_functions()
    assert len(names) >= 10
    for n in names:
        assert hasattr(L, n), f"missing export {n}"
    assert set(_lib.EXPORTS) <= set(names)


def test_struct_sizes_match_library():
    """The library validates struct_bytes; a deliberately wrong size must be rejected with a message."""
    from sat_b200 import _lib
    L = _lib.lib()
    a = _lib.DecoderArgs()
    a.struct_bytes = 4
    n = ctypes.c_int64(0)
    rc = L.sat_decoder_workspace_bytes(ctypes.byref(a), ctypes.byref(n))
    assert rc != 0 and b"size mismatch" in L.sat_last_error()
    a.struct_bytes = ctypes.sizeof(_lib.DecoderArgs)
    a.dtype, a.B, a.L, a.D, a.H, a.V, a.T = 1, 4, 196, 512, 512, 2531, 20
    a.tf = a.attention = a.gate = a.need_grad = 1
    a.alphas = 256
    a.preds = 256
    rc = L.sat_decoder_workspace_bytes(ctypes.byref(a), ctypes.byref(n))
    assert rc == 0, L.sat_last_error()
    assert n.value > 0
    b = _lib.BeamArgs()
    b.struct_bytes = 8
    assert L.sat_beam_workspace_bytes(ctypes.byref(b), ctypes.byref(n)) != 0


def test_product_path_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from sat_b200.decoder import Decoder
    dec = Decoder(50, 16, tf=True, attention=True)
    with pytest.raises(RuntimeError, match="CUDA"):
        dec(torch.zeros(2, 4, 16), torch.zeros(2, 5, dtype=torch.long))
