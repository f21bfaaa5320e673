"""The Tier-2 writer copies the bodies of neighbouring code-blocks as one run when they are contiguous in memory (they are in the
encoder's packed arena).  Same block results once back to back in one arena and once in separate allocations: the codestreams
must be identical, for one and several layers and every progression order."""
import ctypes as C
import numpy as np
import pytest
import hostpipe
from j2k_b200 import _lib as L


@pytest.mark.parametrize("kw", [dict(), dict(numlayers=3, rates=(40, 20, 0)), dict(prog_order=2, tile_w=128, tile_h=96), dict(prog_order=4, numlayers=2, rates=(30, 0)),
                                dict(cblksty=4), dict(cblksty=0x26, numlayers=2, rates=(25, 0))])
def test_contiguous_and_scattered_block_data_give_the_same_stream(kw):
    lib = L.lib()
    p = L.default_params(width=300, height=200, ncomp=3, prec=8, mct=1, **kw)
    n = len(hostpipe.block_list(p))
    rng = np.random.default_rng(n)
    npasses = 10
    rates = [np.cumsum(rng.integers(0, 60, size=npasses)).astype(np.int32) for _ in range(n)]
    dist = [np.cumsum(rng.random(npasses) * 1e5) for _ in range(n)]
    arena = rng.integers(0, 255, size=int(sum(int(r[-1]) for r in rates)) + 16, dtype=np.uint8)

    def assemble(contiguous):
        res = (L.BlockResult * n)()
        keep, off = [], 0
        for i in range(n):
            ln = int(rates[i][-1])
            if contiguous:
                ptr = C.cast(arena.ctypes.data + off, C.POINTER(C.c_uint8))
            else:
                d = arena[off:off + ln + 1].copy(); keep.append(d)
                ptr = d.ctypes.data_as(C.POINTER(C.c_uint8))
            off += ln
            res[i].data = ptr; res[i].numbps = 4; res[i].npasses = npasses
            res[i].rates = rates[i].ctypes.data_as(C.POINTER(C.c_int32)); res[i].dist = dist[i].ctypes.data_as(C.POINTER(C.c_double))
        out = C.POINTER(C.c_uint8)(); ln = C.c_size_t(0)
        assert lib.j2kcuda_t2_assemble(C.byref(p), res, n, C.byref(out), C.byref(ln)) == 0, L.last_error()
        data = C.string_at(out, ln.value)
        lib.j2kcuda_free(out)
        return data

    assert assemble(True) == assemble(False)
