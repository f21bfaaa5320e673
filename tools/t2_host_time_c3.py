"""Time the host Tier-2 writer with the cinema rate cap (PCRD bisection) on synthetic code-block results (no GPU needed).
Prints an md5 of the codestream so that a change of the host code can be checked for identical output."""
import ctypes as C, sys, os, time, hashlib
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from j2k_b200 import _lib as L
lib = L.lib()
p = L.default_params(width=4096, height=2160, ncomp=3, prec=12, irreversible=1, mct=1, profile=L.PROFILE_CINEMA_4K)
nblk = lib.j2kcuda_t2_block_count(C.byref(p))          # the library's own block table (no oracle in tools/)
geo = np.zeros((nblk, 12), dtype=np.int32)
assert lib.j2kcuda_t2_block_geometry(C.byref(p), geo.ctypes.data_as(C.c_void_p), nblk) == nblk
n = len(geo)
rng = np.random.default_rng(0)
results = (L.BlockResult * n)()
keep = []
npasses = int(sys.argv[1]) if len(sys.argv) > 1 else 16
for i in range(n):
    rates = np.cumsum(rng.integers(10, 60, size=npasses)).astype(np.int32)
    data = rng.integers(0, 255, size=int(rates[-1]) + 1, dtype=np.uint8)
    data[data == 255] = 0
    dist = np.cumsum(1e7 * rng.random(npasses) * (0.6 ** np.arange(npasses)))
    keep.append((rates, data, dist))
    results[i].data = data.ctypes.data_as(C.POINTER(C.c_uint8))
    results[i].numbps = (npasses + 2) // 3; results[i].npasses = npasses
    results[i].rates = rates.ctypes.data_as(C.POINTER(C.c_int32))
    results[i].dist = dist.ctypes.data_as(C.POINTER(C.c_double))
for rep in range(3):
    out = C.POINTER(C.c_uint8)(); ln = C.c_size_t(0)
    t0 = time.perf_counter()
    rc = lib.j2kcuda_t2_assemble(C.byref(p), results, n, C.byref(out), C.byref(ln))
    t1 = time.perf_counter()
    assert rc == 0, L.last_error()
    md5 = hashlib.md5(C.string_at(out, ln.value)).hexdigest()
    lib.j2kcuda_free(out)
    print("assemble %d blocks -> %d bytes: %.2f ms  md5 %s" % (n, ln.value, (t1 - t0) * 1e3, md5))
