"""Per-stage device times of the encoder for one configuration (prints the library's CUDA-event timings)."""
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


def run(h, w, nc, prec, kind, reps=5, **kw):
    lib = L.lib()
    img = make_frame(h, w, nc, prec, kind, 0)
    p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, **kw)
    lib.j2kcuda_encoder_create.restype = C.c_void_p
    e = C.c_void_p(lib.j2kcuda_encoder_create(C.byref(p)))
    assert e.value, L.last_error()
    pl = j2k.planes_of(img)
    names = ["upload", "mct", "dwt", "t1ctx", "t1mq", "compact+d2h", "host_t2", "dwt_l1"]
    for r in range(reps):
        t0 = time.perf_counter()
        assert lib.j2kcuda_encoder_upload(e, pl, nc) == 0, L.last_error()
        assert lib.j2kcuda_encoder_run(e) == 0, L.last_error()
        t1 = time.perf_counter()
        out = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
        assert lib.j2kcuda_encoder_finish(e, C.byref(out), C.byref(n)) == 0, L.last_error()
        t2 = time.perf_counter()
        lib.j2kcuda_free(out)
    ms = (C.c_float * 8)(); cnt = (C.c_uint64 * 6)()
    lib.j2kcuda_encoder_timings(e, ms, 8); lib.j2kcuda_encoder_counters(e, cnt, 6)
    print("%dx%dx%d %d-bit %s %s" % (w, h, nc, prec, kind, kw))
    print("  " + "  ".join("%s=%.3f" % (a, b) for a, b in zip(names, ms)))
    print("  wall: upload+run %.2f ms, finish %.2f ms; blocks %d, symbols %d, bytes %d, launches %d" %
          ((t1 - t0) * 1e3, (t2 - t1) * 1e3, cnt[1], cnt[2], cnt[3], cnt[0]))
    npx = w * h
    print("  dwt_l1 %.1f GB/s (algorithmic 8 B/sample)  dwt all %.1f GB/s  MPix/s (run+finish) %.1f" %
          (8.0 * npx * nc / (ms[7] * 1e-3) / 1e9, 8.0 * npx * nc * 1.333 / (ms[2] * 1e-3) / 1e9, npx / (t2 - t0) / 1e6))
    lib.j2kcuda_encoder_destroy(e)


def run_decode(h, w, nc, prec, kind, reps=5, **kw):
    lib = L.lib()
    img = make_frame(h, w, nc, prec, kind, 0)
    cs = j2k.encode(img, prec, **kw)
    lib.j2kcuda_decoder_create.restype = C.c_void_p
    d = C.c_void_p(lib.j2kcuda_decoder_create(-1))
    out = np.zeros_like(img)
    pl = j2k.planes_of(out)
    for r in range(reps):
        t0 = time.perf_counter()
        assert lib.j2kcuda_decoder_prepare(d, cs, C.c_size_t(len(cs)), 0) == 0, L.last_error()
        t1 = time.perf_counter()
        assert lib.j2kcuda_decoder_run(d) == 0, L.last_error()
        t2 = time.perf_counter()
        assert lib.j2kcuda_decoder_download(d, pl, nc) == 0, L.last_error()
        t3 = time.perf_counter()
    assert kw.get("irreversible") or np.array_equal(out, img)
    ms = (C.c_float * 8)()
    lib.j2kcuda_decoder_timings(d, ms, 7)
    names = ["host_parse", "h2d", "t1", "idwt", "imct", "d2h", "idwt_l1"]
    print("DECODE %dx%dx%d %d-bit %s: %d bytes" % (w, h, nc, prec, kind, len(cs)))
    print("  " + "  ".join("%s=%.3f" % (a, b) for a, b in zip(names, ms)))
    print("  wall: prepare %.2f run %.2f download %.2f ms -> %.1f MPix/s" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, w * h / (t3 - t0) / 1e6))
    lib.j2kcuda_decoder_destroy(d)




if __name__ == "__main__":
    if "c3" in sys.argv:
        for kind in ("smooth", "noise"):
            run(2160, 4096, 3, 12, kind, reps=4, irreversible=1, mct=1, profile=L.PROFILE_CINEMA_4K)
            run_decode(2160, 4096, 3, 12, kind, irreversible=1, mct=1, profile=L.PROFILE_CINEMA_4K)
    elif "roundtrip1" in sys.argv:       # the ncu --set full target: every kernel of one 4K RGBA16 encode + decode (one warm pass first)
        run(2160, 4096, 4, 16, "smooth", reps=2, mct=1)
        run_decode(2160, 4096, 4, 16, "smooth", reps=2, mct=1)
    elif "dwt" in sys.argv:
        run(2160, 4096, 4, 16, "smooth", reps=6, mct=1)
    elif "decode1" in sys.argv:
        run_decode(2160, 4096, 4, 16, "smooth", mct=1)
    elif "decode" in sys.argv:
        for kind in ("smooth", "noise"):
            run_decode(2160, 4096, 4, 16, kind, mct=1)
        run_decode(1080, 2048, 3, 8, "smooth", mct=1)
    else:
        for kind in ("smooth", "noise"):
            run(2160, 4096, 4, 16, kind, mct=1)
        run(1080, 2048, 3, 8, "smooth", mct=1)
        for lanes in (1, 4, 8, 16):
            os.environ["J2KCUDA_MQ_LANES"] = str(lanes)
            print("MQ lanes", lanes)
            run(2160, 4096, 4, 16, "smooth", reps=3, mct=1)

