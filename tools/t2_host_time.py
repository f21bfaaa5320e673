"""Time the host Tier-2 writer alone on synthetic code-block results (no GPU needed)."""
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
# the code bytes of all blocks back to back in block order, as the encoder's packed arena holds them after the D2H copy
all_rates = [np.cumsum(rng.integers(20, 80, size=28)).astype(np.int32) for _ in range(n)]
arena = rng.integers(0, 255, size=int(sum(int(r[-1]) for r in all_rates)) + 64, dtype=np.uint8)
off = 0
for i in range(n):
    npasses = 28
    rates = all_rates[i]
    data = arena[off:off + int(rates[-1])]
    off += int(rates[-1])
    dist = np.cumsum(rng.random(npasses) * 1e6)[::1].copy()
    keep.append((rates, data, dist))
    results[i].data = C.cast(arena.ctypes.data + (off - int(rates[-1])), C.POINTER(C.c_uint8))
    results[i].numbps = 10; results[i].npasses = npasses
    results[i].rates = rates.ctypes.data_as(C.POINTER(C.c_int32))
    results[i].dist = dist.ctypes.data_as(C.POINTER(C.c_double))
for rep in range(5):
    out = C.POINTER(C.c_uint8)(); ln = C.c_size_t(0)
    t0 = time.perf_counter()
    rc = lib.j2kcuda_t2_assemble(C.byref(p), results, n, C.byref(out), C.byref(ln))
    t1 = time.perf_counter()
    assert rc == 0, L.last_error()
    lib.j2kcuda_free(out)
    print("assemble %d blocks -> %d bytes: %.2f ms" % (n, ln.value, (t1 - t0) * 1e3))
