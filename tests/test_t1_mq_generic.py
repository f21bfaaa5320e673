"""Row a21 (first part): encoding with the code-block styles RESET and TERMALL.  The generic MQ encoder the CUDA kernel k_t1_mq_gen
runs (j2k_b200/csrc/t1_mq_gen.cuh) compiled for the host, fed the (context, decision) symbols of the oracle's context modelling
(the same sequence k_t1_context writes: tests/test_gpu_stages.py), against the REFERENCE's generic encoder
(opj_t1_encode_cblk, t1.cpp:1537-1673, compiled into oracle/_ref) block by block: bytes and per-pass rates.  The oracle's own
restatement of that loop (orc_t1_encode_cblk_sty) is pinned the same way."""
import ctypes as C
import os
import subprocess
import numpy as np
import pytest
import orc
import refgrok
from test_t1dec_core import Tables

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def core():
    subprocess.check_call(["make", "-C", os.path.join(HERE, "native")], stdout=subprocess.DEVNULL)
    lib = C.CDLL(os.path.join(HERE, "native", "libt1dec_host.so"))
    lib._ZN3j2k14t1_tables_hostEv.restype = C.POINTER(Tables)
    return lib, lib._ZN3j2k14t1_tables_hostEv().contents


def blocks(seed, n):
    rng = np.random.default_rng(seed)
    for it in range(n):
        w, h = int(rng.integers(1, 65)), int(rng.integers(1, 65))
        if it % 5 == 0:
            w, h = 64, 64
        coeff = rng.laplace(0, float(rng.choice([2, 30, 400, 5000])), size=(h, w)).astype(np.int32)
        if it % 19 == 0:
            coeff[:] = 0
        if it % 23 == 1:
            coeff[:] = 0; coeff[h // 2, w // 2] = 12345
        yield coeff, int(rng.integers(0, 4))


@pytest.mark.parametrize("sty", [2, 4, 6])
def test_oracle_generic_loop_equals_the_reference(sty):
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    for coeff, orient in blocks(sty, 60):
        smag, mx = orc.quantize(coeff)
        a = orc.t1_encode_sty(smag, orient, sty, mx)
        b = refgrok.t1_encode_cblk_generic(coeff, orient, sty)
        assert (a["numbps"], a["npasses"], a["rates"]) == (b["numbps"], b["npasses"], b["rates"])
        assert a["data"] == b["data"]
        if sty & 4:
            assert all(b["terms"])


@pytest.mark.parametrize("sty", [2, 4, 6])
def test_host_compiled_device_core_equals_the_reference(core, sty):
    lib, tab = core
    have_ref = refgrok.available()
    for coeff, orient in blocks(100 + sty, 60):
        smag, mx = orc.quantize(coeff)
        o = orc.t1_encode_sty(smag, orient, sty, mx, trace=True)
        k = o["npasses"]
        sym = np.ascontiguousarray(o["symbols"])
        pn = (C.c_uint32 * 128)(*o["pass_nsym"])
        out = np.zeros(coeff.size * 8 + 64, np.uint8)
        rate = (C.c_int32 * 128)()
        total = C.c_int32(0)
        rc = lib.mqgen_host(sym.ctypes.data_as(C.c_void_p), pn, k, sty, tab.mq, out.ctypes.data_as(C.c_void_p), out.size, rate, C.byref(total))
        assert rc == 0
        want = refgrok.t1_encode_cblk_generic(coeff, orient, sty) if have_ref else o
        assert [rate[i] for i in range(k)] == want["rates"]
        assert out[:total.value].tobytes() == want["data"]


def test_core_reports_a_full_slot_instead_of_writing_past_it(core):
    lib, tab = core
    coeff = np.random.default_rng(5).laplace(0, 3000, size=(64, 64)).astype(np.int32)
    smag, mx = orc.quantize(coeff)
    o = orc.t1_encode_sty(smag, 1, 4, mx, trace=True)
    sym = np.ascontiguousarray(o["symbols"])
    pn = (C.c_uint32 * 128)(*o["pass_nsym"])
    out = np.full(256 + 8, 0xAB, np.uint8)
    rate = (C.c_int32 * 128)(); total = C.c_int32(0)
    rc = lib.mqgen_host(sym.ctypes.data_as(C.c_void_p), pn, o["npasses"], 4, tab.mq, out.ctypes.data_as(C.c_void_p), 256, rate, C.byref(total))
    assert rc == 2 and total.value == 0
    assert (out[256:] == 0xAB).all()
