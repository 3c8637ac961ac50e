"""The C-ABI library loads here (no GPU) and exports every symbol include/prt.h declares; no compute
is called. Also: the product fails loudly, without any CPU fallback, when there is no device."""
import ctypes
import os
import re

import pytest

from parallelraytracing_b200 import build as B
from parallelraytracing_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "prt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(prt_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(capi.EXPORTS)


def test_library_exports_every_declared_symbol():
    so = B.build()
    lib = ctypes.CDLL(so)
    for name in declared_symbols():
        assert hasattr(lib, name), name
    lib.prt_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.prt_version()


def test_struct_sizes_match_header():
    assert ctypes.sizeof(capi.CCamera) == 4 * 13 + 8
    assert capi.sphere_dt.itemsize == 32 and capi.quad_dt.itemsize == 80 and capi.material_dt.itemsize == 32
    assert capi.hit_dt.itemsize == 48 and capi.ray_dt.itemsize == 24
    assert ctypes.sizeof(capi.CStats) == 8 + 8 * 32 + 8 * 3 + 4 * 6 + 4 * 8


def test_fails_loudly_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.PrtError):
        capi.Context(0)


def test_library_is_sm100a_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", capi.lib_path()], capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out and "sm_80" not in out
