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
