"""Sequences (config C5: frames of an image sequence on one GPU): j2kcuda_sequence_encode / _decode keep several
frames in flight on worker threads; every frame must come out byte-identical to the one-frame call and decode
bit-exactly, in host-pixel and device-pixel mode, and errors of any worker must surface."""
import ctypes as C
import numpy as np
import pytest
import j2k
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame

pytestmark = pytest.mark.gpu


def _table(arrays):
    keep = [j2k.planes_of(a) for a in arrays]
    return (C.POINTER(L.Plane) * len(keep))(*[C.cast(k, C.POINTER(L.Plane)) for k in keep]), keep


@pytest.mark.parametrize("workers", [1, 3])
@pytest.mark.parametrize("h,w,nc,prec", [(270, 480, 3, 8), (200, 340, 4, 16)])
def test_sequence_matches_single_frame_calls(workers, h, w, nc, prec):
    lib = L.lib()
    lib.j2kcuda_sequence_create.restype = C.c_void_p
    n = 7
    frames = [make_frame(h, w, nc, prec, ("smooth", "noise", "ramp")[i % 3], seed=i) for i in range(n)]
    p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, mct=1)
    seq = C.c_void_p(lib.j2kcuda_sequence_create(C.byref(p), -1, workers))
    assert seq.value
    tin, _k = _table(frames)
    outs = (C.POINTER(C.c_uint8) * n)()
    lens = (C.c_size_t * n)()
    for rep in range(2):                      # the contexts are reused by the second call
        assert lib.j2kcuda_sequence_encode(seq, tin, n, nc, 0, outs, lens) == 0, L.last_error()
        streams = [C.string_at(outs[i], lens[i]) for i in range(n)]
        for i in range(n):
            assert streams[i] == j2k.encode(frames[i], prec, mct=1), "frame %d differs from the single-frame encode" % i
        back = [np.zeros_like(f) for f in frames]
        tout, _k2 = _table(back)
        assert lib.j2kcuda_sequence_decode(seq, C.cast(outs, C.POINTER(C.POINTER(C.c_uint8))), lens, n, 0, tout, nc, 0) == 0, L.last_error()
        for i in range(n):
            assert np.array_equal(back[i], frames[i])
            lib.j2kcuda_free(outs[i])
    fr = (C.c_uint64 * 2)()
    assert lib.j2kcuda_sequence_stats(seq, None, None, None, None, fr, 1) == 0
    assert list(fr) == [2 * n, 2 * n]
    lib.j2kcuda_sequence_destroy(seq)


def test_sequence_device_pixels_and_one_shot_batch():
    import torch
    lib = L.lib()
    lib.j2kcuda_sequence_create.restype = C.c_void_p
    h, w, nc, prec, n = 135, 240, 4, 16, 5
    frames = [make_frame(h, w, nc, prec, "smooth", seed=10 + i) for i in range(n)]
    dev = [torch.from_numpy(f.view(np.int16)).cuda() for f in frames]
    out = [torch.zeros_like(d) for d in dev]

    def table(ts):
        keep = []
        for t in ts:
            pl = (L.Plane * 1)()
            pl[0].buf = t.data_ptr(); pl[0].colbytes = nc * 2; pl[0].rowbytes = w * nc * 2
            pl[0].width = w; pl[0].height = h; pl[0].sample_type = L.USHORT
            keep.append(pl)
        return (C.POINTER(L.Plane) * n)(*[C.cast(k, C.POINTER(L.Plane)) for k in keep]), keep
    tin, _a = table(dev)
    tout, _b = table(out)
    p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, mct=1)
    seq = C.c_void_p(lib.j2kcuda_sequence_create(C.byref(p), -1, 2))
    outs = (C.POINTER(C.c_uint8) * n)()
    lens = (C.c_size_t * n)()
    assert lib.j2kcuda_sequence_encode(seq, tin, n, nc, 1, outs, lens) == 0, L.last_error()
    for i in range(n):
        assert C.string_at(outs[i], lens[i]) == j2k.encode(frames[i], prec, mct=1)
    assert lib.j2kcuda_sequence_decode(seq, C.cast(outs, C.POINTER(C.POINTER(C.c_uint8))), lens, n, 0, tout, nc, 1) == 0, L.last_error()
    torch.cuda.synchronize()
    for i in range(n):
        assert torch.equal(out[i], dev[i])
    lib.j2kcuda_sequence_destroy(seq)
    # one-shot batch forms with host pixels
    hin, _c = _table(frames)
    back = [np.zeros_like(f) for f in frames]
    hout, _d = _table(back)
    outs2 = (C.POINTER(C.c_uint8) * n)()
    lens2 = (C.c_size_t * n)()
    assert lib.j2kcuda_encode_batch(C.byref(p), hin, n, nc, outs2, lens2) == 0, L.last_error()
    for i in range(n):
        assert C.string_at(outs2[i], lens2[i]) == C.string_at(outs[i], lens[i])
    assert lib.j2kcuda_decode_batch(C.cast(outs2, C.POINTER(C.POINTER(C.c_uint8))), lens2, n, 0, hout, nc) == 0, L.last_error()
    for i in range(n):
        assert np.array_equal(back[i], frames[i])
        lib.j2kcuda_free(outs[i]); lib.j2kcuda_free(outs2[i])


def test_sequence_reports_worker_errors():
    lib = L.lib()
    lib.j2kcuda_sequence_create.restype = C.c_void_p
    seq = C.c_void_p(lib.j2kcuda_sequence_create(None, -1, 3))
    good = j2k.encode(make_frame(64, 64, 1, 8, "ramp", 0), 8)
    bad = b"\xff\x4f\xff\x51" + b"\x00" * 40
    streams = [good, bad, good, good]
    n = len(streams)
    bufs = [C.create_string_buffer(s, len(s)) for s in streams]
    ptrs = (C.POINTER(C.c_uint8) * n)(*[C.cast(b, C.POINTER(C.c_uint8)) for b in bufs])
    lens = (C.c_size_t * n)(*[len(s) for s in streams])
    back = [np.zeros((64, 64, 1), np.uint8) for _ in range(n)]
    tout, _k = _table(back)
    rc = lib.j2kcuda_sequence_decode(seq, ptrs, lens, n, 0, tout, 1, 0)
    assert rc == L.ERR_STREAM, (rc, L.last_error())
    assert L.last_error()
    lib.j2kcuda_sequence_destroy(seq)


def test_one_shot_calls_are_reentrant():
    """The plug-in's codec objects are process-wide singletons and After Effects calls them from several render threads
    (threading): concurrent j2kcuda_encode / j2kcuda_decode calls with different parameters must not disturb
    each other.  Every thread's codestream must equal the one the same call gives alone."""
    import threading
    jobs = [(96, 128, 3, 8, dict(mct=1)), (130, 200, 4, 16, dict(mct=1)), (75, 133, 1, 8, dict(numresolution=4, cblk_w=32, cblk_h=32)),
            (100, 160, 3, 12, dict(irreversible=1, mct=1, rates=[10.0])), (200, 340, 3, 16, dict(mct=1, tile_w=64, tile_h=64)),
            (64, 64, 3, 8, dict(mct=1, prog_order=3))]
    imgs = [make_frame(h, w, nc, prec, "smooth" if i % 2 else "noise", seed=i) for i, (h, w, nc, prec, kw) in enumerate(jobs)]
    alone = [j2k.encode(img, job[3], **job[4]) for img, job in zip(imgs, jobs)]
    decoded_alone = [j2k.decode(cs) for cs in alone]
    errors = []

    def worker(i):
        try:
            for _ in range(4):
                cs = j2k.encode(imgs[i], jobs[i][3], **jobs[i][4])
                if cs != alone[i]:
                    errors.append("thread %d: codestream differs" % i)
                if not np.array_equal(j2k.decode(cs), decoded_alone[i]):
                    errors.append("thread %d: decoded pixels differ" % i)
        except AssertionError as e:
            errors.append("thread %d: %s" % (i, e))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(len(jobs))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


# ---- several GPUs behind one call (here: the same device listed more than once, which exercises the same code paths) ----
def _devices(n):
    import torch
    have = torch.cuda.device_count()
    return [i % have for i in range(n)]


@pytest.mark.gpu
@pytest.mark.parametrize("ndev", [2, 3])
def test_tiles_of_one_frame_over_several_devices(ndev):
    """j2kcuda_tiles_*: contiguous tile runs per device, fragments joined on the host: byte-identical to a one-device encode;
    decode routes tile-parts by SOT tile index and every device writes its own rectangles"""
    import j2k
    lib = L.lib()
    img = make_frame(300, 420, 3, 16, "smooth", seed=21)
    kw = dict(mct=1, tile_w=128, tile_h=128, numresolution=4)
    whole = j2k.encode(img, 16, **kw)
    p = L.default_params(width=420, height=300, ncomp=3, prec=16, **kw)
    devs = (C.c_int * ndev)(*_devices(ndev))
    lib.j2kcuda_tiles_create.restype = C.c_void_p
    t = C.c_void_p(lib.j2kcuda_tiles_create(C.byref(p), devs, ndev))
    assert t.value, L.last_error()
    for _ in range(2):                                               # second frame reuses the contexts
        out = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
        assert lib.j2kcuda_tiles_encode(t, j2k.planes_of(img), 3, C.byref(out), C.byref(n)) == 0, L.last_error()
        data = C.string_at(out, n.value); lib.j2kcuda_free(out)
        assert data == whole
        back = np.zeros_like(img)
        assert lib.j2kcuda_tiles_decode(t, data, C.c_size_t(len(data)), 0, j2k.planes_of(back), 3) == 0, L.last_error()
        assert np.array_equal(back, img)
    half = np.zeros((150, 210, 3), np.uint16)
    assert lib.j2kcuda_tiles_decode(t, whole, C.c_size_t(len(whole)), 1, j2k.planes_of(half), 3) == 0, L.last_error()
    assert np.array_equal(half, j2k.decode(whole, reduce=1))
    lib.j2kcuda_tiles_destroy(t)


@pytest.mark.gpu
def test_sequence_over_several_devices():
    import j2k
    lib = L.lib()
    frames = [make_frame(96, 160, 3, 8, "smooth", seed=s) for s in range(7)]
    p = L.default_params(width=160, height=96, ncomp=3, prec=8, mct=1)
    devs = (C.c_int * 2)(*_devices(2))
    lib.j2kcuda_sequence_create_multi.restype = C.c_void_p
    s = C.c_void_p(lib.j2kcuda_sequence_create_multi(C.byref(p), devs, 2, 2))
    assert s.value, L.last_error()
    keep = [j2k.planes_of(f) for f in frames]
    tab = (C.POINTER(L.Plane) * len(frames))(*[C.cast(k, C.POINTER(L.Plane)) for k in keep])
    outs = (C.POINTER(C.c_uint8) * len(frames))(); lens = (C.c_size_t * len(frames))()
    assert lib.j2kcuda_sequence_encode(s, tab, len(frames), 3, 0, outs, lens) == 0, L.last_error()
    for i, f in enumerate(frames):
        assert C.string_at(outs[i], lens[i]) == j2k.encode(f, 8, mct=1)
    back = [np.zeros_like(f) for f in frames]
    keep2 = [j2k.planes_of(b) for b in back]
    tab2 = (C.POINTER(L.Plane) * len(frames))(*[C.cast(k, C.POINTER(L.Plane)) for k in keep2])
    assert lib.j2kcuda_sequence_decode(s, C.cast(outs, C.POINTER(C.POINTER(C.c_uint8))), lens, len(frames), 0, tab2, 3, 0) == 0, L.last_error()
    for b, f in zip(back, frames):
        assert np.array_equal(b, f)
    for i in range(len(frames)):
        lib.j2kcuda_free(outs[i])
    lib.j2kcuda_sequence_destroy(s)
