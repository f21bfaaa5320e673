"""Damaged codestreams through the WHOLE decoder on the GPU (the round-1 fuzz stopped at the host parser): random damage inside
packet headers (inclusion / Lblock / length fields) and inside code-block bodies.  Whatever the bytes, j2kcuda_decode returns a
status or pixels -- never a fault: a decode of the intact stream afterwards still works on the same CUDA context."""
import ctypes as C
import numpy as np
import pytest
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame

pytestmark = pytest.mark.gpu
ALLOWED = (L.OK, L.ERR_STREAM, L.ERR_UNSUPPORTED, L.ERR_PARAM, L.ERR_NOMEM)


def _decode_status(data, shape, dtype):
    import j2k
    img = np.zeros(shape, dtype=dtype)
    pl = j2k.planes_of(img)
    return L.lib().j2kcuda_decode(data, C.c_size_t(len(data)), 0, pl, shape[2], None, None), img


@pytest.mark.parametrize("kw,prec,kind", [(dict(mct=1), 8, "smooth"), (dict(mct=1, cblk_w=32, cblk_h=32, numlayers=3, rates=(30, 10, 0)), 16, "noise"),
                                          (dict(irreversible=1, mct=1, tile_w=64, tile_h=64, numresolution=4), 12, "smooth")])
def test_damage_in_packet_headers_and_bodies_never_faults(kw, prec, kind):
    import j2k
    dtype = np.uint8 if prec <= 8 else np.uint16
    img = make_frame(96, 128, 3, prec, kind, seed=prec)
    cs = j2k.encode(img, prec, **kw)
    sod = cs.find(b"\xff\x93") + 2
    assert sod > 2
    rng = np.random.default_rng(prec)
    seen = {}
    for it in range(170):
        bad = bytearray(cs)
        mode = it % 4
        if mode == 0:                                  # a few random bytes anywhere after the first SOD
            for k in rng.integers(sod, len(cs) - 2, size=int(rng.integers(1, 6))):
                bad[int(k)] = int(rng.integers(0, 256))
        elif mode == 1:                                # the first packet headers: inclusion tag trees, pass counts, Lblock, lengths
            for k in rng.integers(sod, min(sod + 48, len(cs) - 2), size=3):
                bad[int(k)] ^= 1 << int(rng.integers(0, 8))
        elif mode == 2:                                # a run of 0xFF (markers / bit stuffing) inside a body
            k = int(rng.integers(sod, len(cs) - 40))
            bad[k:k + int(rng.integers(2, 24))] = b"\xff" * 24
        else:                                          # truncation + garbage tail
            k = int(rng.integers(sod, len(cs)))
            bad = bad[:k] + bytearray(rng.integers(0, 256, size=int(rng.integers(0, 64)), dtype=np.uint8).tobytes())
        rc, _ = _decode_status(bytes(bad), img.shape, dtype)
        assert rc in ALLOWED, (it, rc, L.last_error())
        assert rc != L.ERR_CUDA
        seen[rc] = seen.get(rc, 0) + 1
    rc, out = _decode_status(cs, img.shape, dtype)     # the context survived: the intact stream still decodes
    assert rc == 0, L.last_error()
    if not kw.get("irreversible") and not kw.get("numlayers"):
        assert np.array_equal(out, img)
    assert seen.get(L.OK, 0) > 0                       # damage inside bodies decodes to (wrong) pixels, as in the reference
