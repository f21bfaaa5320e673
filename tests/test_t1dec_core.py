"""The Tier-1 decoder block routine the CUDA kernel runs (j2k_b200/csrc/t1_dec_core.cuh), compiled for the
host with one lane, against the oracle: every truncation point of several blocks.  (No GPU needed; the GPU
run of the same routine is tests/test_gpu_decode.py.)"""
import ctypes as C
import os
import subprocess
import numpy as np
import pytest
import orc

HERE = os.path.dirname(os.path.abspath(__file__))


class Tables(C.Structure):
    _fields_ = [("zc", C.c_uint8 * 1024), ("sc", C.c_uint8 * 256), ("nm", C.c_int16 * 512), ("mq", C.c_uint32 * 94),
                ("pad", C.c_uint32 * 2), ("zc9", C.c_uint8 * 2048)]


@pytest.fixture(scope="module")
def core():
    subprocess.check_call(["make", "-C", os.path.join(HERE, "native")], stdout=subprocess.DEVNULL)
    lib = C.CDLL(os.path.join(HERE, "native", "libt1dec_host.so"))
    lib._ZN3j2k14t1_tables_hostEv.restype = C.POINTER(Tables)
    return lib, lib._ZN3j2k14t1_tables_hostEv().contents


def run(core, data, w, h, orient, numbps, npasses):
    lib, tab = core
    out = np.zeros((h, w), np.int32)
    buf = np.frombuffer(data + b"\0", dtype=np.uint8).copy()
    lib.t1dec_host_block(buf.ctypes.data_as(C.c_void_p), len(data), w, h, orient, numbps, npasses, tab.zc9, tab.sc, tab.mq,
                         out.ctypes.data_as(C.c_void_p))
    return out


@pytest.mark.parametrize("w,h,scale", [(4, 4, 5), (8, 8, 40), (64, 64, 40), (17, 9, 3000), (64, 64, 3), (32, 32, 20000),
                                       (64, 3, 300), (5, 64, 300), (1, 1, 9), (33, 37, 150)])
@pytest.mark.parametrize("orient", [0, 1, 3])
def test_every_truncation_point(core, w, h, scale, orient):
    rng = np.random.default_rng(w * 1000 + h + scale)
    c = rng.laplace(0, scale, size=(h, w)).astype(np.int32)
    if scale <= 5:
        c[rng.random((h, w)) < 0.6] = 0
    q, mx = orc.quantize(c)
    r = orc.t1_encode(q, orient, mx)
    for npasses in range(1, r["npasses"] + 1):
        data = r["data"][:r["rates"][npasses - 1]]
        want = orc.t1_decode(data, w, h, orient, r["numbps"], npasses)
        got = run(core, data, w, h, orient, r["numbps"], npasses)
        assert np.array_equal(got, want), "npasses %d of %d" % (npasses, r["npasses"])
    if r["npasses"]:
        assert np.array_equal(np.fix(got / 2).astype(np.int32), c)


# ---- code-block styles (row f4): the generic variant of the same routine against the reference's generic encoder and decoder ----
import refgrok

LAZY, RESET, TERMALL, VSC, PTERM, SEGSYM = 1, 2, 4, 8, 16, 32


def segments_of(r, sty):
    """codeword segments of a block from the per-pass term flags of the reference encoder (t1.cpp:1614-1628)"""
    segs, start, npass = [], 0, 0
    for i in range(r["npasses"]):
        npass += 1
        if r["terms"][i] or i == r["npasses"] - 1:
            segs.append((r["rates"][i] - start, npass))
            start, npass = r["rates"][i], 0
    return segs


@pytest.mark.skipif(not refgrok.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("sty", [RESET, TERMALL, VSC, SEGSYM, LAZY, PTERM, LAZY | TERMALL, RESET | VSC | SEGSYM, LAZY | RESET | TERMALL | VSC | PTERM | SEGSYM])
@pytest.mark.parametrize("w,h,scale,orient", [(64, 64, 40, 0), (32, 32, 3000, 1), (17, 9, 300, 3), (64, 3, 20, 2), (5, 64, 5, 1), (33, 37, 20000, 0)])
def test_code_block_styles_against_the_reference_decoder(core, sty, w, h, scale, orient):
    lib, tab = core
    rng = np.random.default_rng(w * 31 + h + scale + sty)
    c = rng.laplace(0, scale, size=(h, w)).astype(np.int32)
    r = refgrok.t1_encode_cblk_generic(c, orient, sty)
    if not r["npasses"]:
        return
    segs = segments_of(r, sty)
    assert sum(s[0] for s in segs) == len(r["data"]) and sum(s[1] for s in segs) == r["npasses"]
    for keep in sorted({len(segs), max(1, len(segs) // 2), 1}):           # all segments, and truncations at segment boundaries
        sub = segs[:keep]
        nbytes = sum(s[0] for s in sub)
        want = refgrok.t1_decode_cblk(r["data"][:nbytes], w, h, orient, r["numbps"], sub, cblksty=sty)
        out = np.zeros((h, w), np.int32)
        buf = np.frombuffer(r["data"][:nbytes] + b"\0", dtype=np.uint8).copy()
        sl = (C.c_int * keep)(*[s[0] for s in sub]); sp = (C.c_int * keep)(*[s[1] for s in sub])
        lib.t1dec_host_block_gen(buf.ctypes.data_as(C.c_void_p), sl, sp, keep, w, h, orient, r["numbps"], sty, 0, tab.zc9, tab.sc, tab.mq,
                                 out.ctypes.data_as(C.c_void_p))
        assert np.array_equal(out, want), "style %d, %d of %d segments" % (sty, keep, len(segs))
    if not (sty & (LAZY | TERMALL)):                                  # one segment: the reference decoder gives the input back
        assert np.array_equal(np.fix(want / 2).astype(np.int32), c)
