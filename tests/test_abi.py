"""The C-ABI library loads on a machine without a GPU, exports every symbol include/j2kcuda.h declares,
and refuses pixel work with J2KCUDA_ERR_NO_DEVICE instead of falling back to a CPU path."""
import ctypes as C
import os
import re
import numpy as np
import pytest
from j2k_b200 import _lib as L

HDR = os.path.join(os.path.dirname(__file__), "..", "include", "j2kcuda.h")


def declared():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(j2kcuda_[a-z0-9_]+)\s*\(", src)) - {"j2kcuda_abort_fn", "j2kcuda_log_fn"})


def test_exports_every_declared_symbol():
    lib = L.lib()
    names = declared()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_struct_sizes_match_header():
    # sizes printed by a C program compiled against include/j2kcuda.h (gcc, x86-64)
    assert (C.sizeof(L.Plane), C.sizeof(L.Params), C.sizeof(L.Info)) == (40, 232, 144)
    assert (C.sizeof(L.Jp2Info), C.sizeof(L.Jp2Params)) == (400, 168)
    p = L.default_params()
    assert (p.numresolution, p.cblk_w, p.cblk_h, p.numlayers, p.device) == (6, 64, 64, 1, -1)


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = L.lib()
    assert lib.j2kcuda_device_count() == 0
    a = np.zeros(16, dtype=np.int32)
    rc = lib.j2kcuda_stage_mct(a.ctypes.data_as(C.c_void_p), a.ctypes.data_as(C.c_void_p),
                               a.ctypes.data_as(C.c_void_p), C.c_size_t(16), 0, 0)
    assert rc == L.ERR_NO_DEVICE
    assert "no CPU path" in L.last_error()
    p = L.default_params(width=16, height=16, ncomp=1, prec=8)
    assert not lib.j2kcuda_encoder_create(C.byref(p))
