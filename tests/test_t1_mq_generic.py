"""Row a21: encoding with the code-block styles RESET, TERMALL, PTERM and SEGSYM (the ones the arithmetic-coder stage can do alone).  The generic MQ encoder the CUDA kernel k_t1_mq_gen
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


REF_DEFECT = 0x100     # orc_t1_encode_cblk_sty: keep the reference's zero length for a terminated pass of exactly one byte


def seg_lens(r):
    return [r["rates"][0]] + [r["rates"][i] - r["rates"][i - 1] for i in range(1, r["npasses"])]


def decodes_back(r, coeff, orient, sty):
    """through the REFERENCE's Tier-1 decoder (opj_t1_decode_cblk), one segment per pass under TERMALL"""
    if r["npasses"] == 0:
        return not coeff.any()
    h, w = coeff.shape
    segs = [(l, 1) for l in seg_lens(r)] if sty & 4 else [(r["rates"][-1], r["npasses"])]
    out = refgrok.t1_decode_cblk(r["data"], w, h, orient, r["numbps"], segs, cblksty=sty)
    o = out.astype(np.int64)                       # the decoder's representation carries one extra fractional bit (t1.cpp:1337-1448)
    return np.array_equal(np.sign(o) * (np.abs(o) >> 1), coeff.astype(np.int64))


STYLES = [0x02, 0x04, 0x06, 0x10, 0x20, 0x30, 0x14, 0x24, 0x36]     # RESET 2, TERMALL 4, PTERM 0x10, SEGSYM 0x20 and combinations


@pytest.mark.parametrize("sty", STYLES)
def test_oracle_generic_loop_equals_the_reference(sty):
    """the restatement with the reference's defect kept is the reference, byte for byte and rate for rate; without it the two
    differ exactly in the blocks the reference cannot read back itself"""
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    defects = 0
    for coeff, orient in blocks(sty, 80):
        smag, mx = orc.quantize(coeff)
        b = refgrok.t1_encode_cblk_generic(coeff, orient, sty)
        a = orc.t1_encode_sty(smag, orient, sty | REF_DEFECT, mx)
        assert (a["numbps"], a["npasses"], a["rates"]) == (b["numbps"], b["npasses"], b["rates"])
        assert a["data"] == b["data"]
        if sty & 4:
            assert all(b["terms"])
        c = orc.t1_encode_sty(smag, orient, sty, mx)
        if sty & 0x10:
            continue                      # PTERM: the restatement is pinned, nothing more (the reference's own blocks do not all decode)
        assert decodes_back(c, coeff, orient, sty)
        if c["rates"] != b["rates"]:
            # t1.cpp:1649-1651: numbytes == 0 after a termination is "one byte", the reference says "none"
            assert sty & 4 and 0 in seg_lens(b) and not decodes_back(b, coeff, orient, sty)
            defects += 1
        else:
            assert c["data"] == b["data"]
    assert sty not in (4, 6) or defects > 0     # the small blocks of these samples do hit it


@pytest.mark.parametrize("sty", [s for s in STYLES if not s & 0x10])         # the device core has no PTERM (see t1_mq_gen.cuh)
def test_host_compiled_device_core_equals_the_oracle_and_decodes_in_the_reference(core, sty):
    lib, tab = core
    have_ref = refgrok.available()
    for coeff, orient in blocks(100 + sty, 80):
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
        got = dict(data=out[:total.value].tobytes(), rates=[rate[i] for i in range(k)], npasses=k, numbps=o["numbps"])
        assert got["rates"] == o["rates"] and got["data"] == o["data"]
        if have_ref:
            assert decodes_back(got, coeff, orient, sty)


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
