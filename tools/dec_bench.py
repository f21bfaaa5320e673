"""Decoder experiment: one 4K RGBA16 frame alone (per-stage CUDA-event times) and a batch of distinct frames through
j2kcuda_sequence_decode with several contexts in flight (MPix/s, device-resident output).  The Tier-1 decoder's shape is
chosen by the environment (J2KCUDA_DEC_IMPL / _LPB / _WARPS / _CTAS, read once per process)."""
import argparse
import ctypes as C
import os
import sys
import time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame
import j2k

ap = argparse.ArgumentParser()
ap.add_argument("--kind", default="smooth")
ap.add_argument("--frames", type=int, default=16)
ap.add_argument("--workers", type=int, default=16)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--shape", default="2160,4096,4,16")
ap.add_argument("--tag", default="")
ap.add_argument("--alone-only", action="store_true")
a = ap.parse_args()
H, W, NC, PREC = [int(v) for v in a.shape.split(",")]
lib = L.lib()
imgs = [make_frame(H, W, NC, PREC, a.kind, s) for s in range(a.frames)]
css = [j2k.encode(im, PREC, mct=1 if NC >= 3 else 0) for im in imgs]

# one frame alone
d = C.c_void_p(lib.j2kcuda_decoder_create(-1))
out = np.zeros_like(imgs[0]); pl = j2k.planes_of(out)
best = None
for r in range(4):
    assert lib.j2kcuda_decoder_prepare(d, css[0], C.c_size_t(len(css[0])), 0) == 0, L.last_error()
    assert lib.j2kcuda_decoder_run(d) == 0, L.last_error()
    assert lib.j2kcuda_decoder_download(d, pl, NC) == 0, L.last_error()
    ms = (C.c_float * 8)(); lib.j2kcuda_decoder_timings(d, ms, 7)
    if r and (best is None or ms[2] < best[2]):
        best = list(ms)
ok_alone = bool(np.array_equal(out, imgs[0]))
lib.j2kcuda_decoder_destroy(d)
if a.alone_only:
    print("%s alone: t1=%.3f ms lossless=%s" % (a.tag, best[2], ok_alone))
    sys.exit(0)

# batch
p = L.default_params(width=W, height=H, ncomp=NC, prec=PREC, mct=1 if NC >= 3 else 0)
lib.j2kcuda_sequence_create.restype = C.c_void_p
seq = C.c_void_p(lib.j2kcuda_sequence_create(C.byref(p), 0, a.workers))
F = a.frames
bufs = [np.frombuffer(c, dtype=np.uint8).copy() for c in css]
ptrs = (C.POINTER(C.c_uint8) * F)(*[b.ctypes.data_as(C.POINTER(C.c_uint8)) for b in bufs])
lens = (C.c_size_t * F)(*[len(c) for c in css])
outs = [np.zeros_like(imgs[0]) for _ in range(F)]
keep = [j2k.planes_of(o) for o in outs]
tab = (C.POINTER(L.Plane) * F)(*[C.cast(k, C.POINTER(L.Plane)) for k in keep])
times = []
for r in range(a.reps + 1):
    t0 = time.perf_counter()
    rc = lib.j2kcuda_sequence_decode(seq, ptrs, lens, F, 0, tab, NC, 0)
    assert rc == 0, L.last_error()
    times.append(time.perf_counter() - t0)
ok_batch = all(np.array_equal(o, i) for o, i in zip(outs, imgs))
dm = (C.c_double * 8)(); dc = (C.c_uint64 * 6)(); em = (C.c_double * 8)(); ec = (C.c_uint64 * 6)(); fr = (C.c_uint64 * 2)()
lib.j2kcuda_sequence_stats(seq, em, ec, dm, dc, fr, 1)
lib.j2kcuda_sequence_destroy(seq)
tb = min(times[1:])
print("%s kind=%s impl=%s lpb=%s warps=%s ctas=%s | alone: t1=%.3f ms parse=%.2f idwt=%.3f lossless=%s | batch %d frames x %d workers: %.1f MPix/s (host out, pageable) t1/frame in batch %.2f ms lossless=%s" % (
    a.tag, a.kind, os.environ.get("J2KCUDA_DEC_IMPL", "1"), os.environ.get("J2KCUDA_DEC_LPB", "4"), os.environ.get("J2KCUDA_DEC_WARPS", "3"),
    os.environ.get("J2KCUDA_DEC_CTAS", "0"), best[2], best[0], best[3], ok_alone, F, a.workers, F * W * H / tb / 1e6, dm[2] / max(fr[1], 1), ok_batch), flush=True)
