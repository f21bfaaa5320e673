"""Time the host Tier-2 parser alone on a 4K RGBA16-sized codestream with synthetic code-block bodies (no GPU needed)."""
import ctypes as C, sys, os, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from j2k_b200 import _lib as L
lib = L.lib()
p = L.default_params(width=4096, height=2160, ncomp=4, prec=16, mct=1, numresolution=6)
nblk = lib.j2kcuda_t2_block_count(C.byref(p))          # the library's own block table (no oracle in tools/)
geo = np.zeros((nblk, 12), dtype=np.int32)
assert lib.j2kcuda_t2_block_geometry(C.byref(p), geo.ctypes.data_as(C.c_void_p), nblk) == nblk
n = len(geo)
rng = np.random.default_rng(0)
results = (L.BlockResult * n)()
keep = []
for i in range(n):
    npasses = 28
    rates = np.cumsum(rng.integers(20, 80, size=npasses)).astype(np.int32)
    data = rng.integers(0, 254, size=int(rates[-1]) + 1, dtype=np.uint8)      # no 0xFF: no marker look-alikes
    dist = np.cumsum(rng.random(npasses) * 1e6).copy()
    keep.append((rates, data, dist))
    results[i].data = data.ctypes.data_as(C.POINTER(C.c_uint8))
    results[i].numbps = 10; results[i].npasses = npasses
    results[i].rates = rates.ctypes.data_as(C.POINTER(C.c_int32))
    results[i].dist = dist.ctypes.data_as(C.POINTER(C.c_double))
out = C.POINTER(C.c_uint8)(); ln = C.c_size_t(0)
assert lib.j2kcuda_t2_assemble(C.byref(p), results, n, C.byref(out), C.byref(ln)) == 0, L.last_error()
cs = C.string_at(out, ln.value)
for rep in range(6):
    info = C.POINTER(C.c_int32)(); nb = C.c_int(0); seg = C.POINTER(C.c_uint8)(); sl = C.c_size_t(0)
    t0 = time.perf_counter()
    rc = lib.j2kcuda_t2_parse(cs, C.c_size_t(len(cs)), 0, C.byref(info), C.byref(nb), C.byref(seg), C.byref(sl))
    t1 = time.perf_counter()
    assert rc == 0, L.last_error()
    print("parse %d bytes -> %d blocks, %d segment bytes: %.2f ms" % (len(cs), nb.value, sl.value, (t1 - t0) * 1e3))
    lib.j2kcuda_free(C.cast(info, C.c_void_p)); lib.j2kcuda_free(C.cast(seg, C.c_void_p))
