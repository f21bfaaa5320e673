"""ctypes wrapper over oracle/_ref/libgrokref.so (the reference's ext/grok compiled from
/root/reference by oracle/Makefile).  Test infrastructure only."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "..", "oracle", "_ref", "libgrokref.so")


class RefParams(C.Structure):
    _fields_ = [("irreversible", C.c_int32), ("numresolution", C.c_int32),
                ("cblockw", C.c_int32), ("cblockh", C.c_int32), ("tcp_mct", C.c_int32),
                ("prog_order", C.c_int32), ("tile_w", C.c_int32), ("tile_h", C.c_int32),
                ("numlayers", C.c_int32), ("disto_alloc", C.c_int32), ("rsiz", C.c_int32),
                ("max_comp_size", C.c_int32), ("max_cs_size", C.c_int64),
                ("num_threads", C.c_int32), ("rates", C.c_double * 8)]


_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_PATH)
        _lib.ref_dwt_getnorm.restype = C.c_double
        _lib.ref_mct_getnorm.restype = C.c_double
        _lib.ref_version.restype = C.c_char_p
    return _lib


def _planes_ptr(planes):
    arr = (C.c_void_p * len(planes))()
    for i, p in enumerate(planes):
        arr[i] = p.ctypes.data
    return arr


def encode(planes, prec, sgnd=0, irreversible=0, numresolution=6, cblk=(64, 64), mct=0, prog=0,
           tile=(0, 0), numlayers=1, disto_alloc=1, rsiz=0, max_comp_size=0, max_cs_size=0,
           threads=1, rates=()):
    """planes: list of (h,w) int32 arrays -> bytes"""
    planes = [np.ascontiguousarray(p, dtype=np.int32) for p in planes]
    h, w = planes[0].shape
    p = RefParams(irreversible, numresolution, cblk[0], cblk[1], mct, prog, tile[0], tile[1],
                  numlayers, disto_alloc, rsiz, max_comp_size, max_cs_size, threads)
    for i, r in enumerate(rates):
        p.rates[i] = r
    cap = w * h * len(planes) * 4 + (1 << 20)
    out = np.empty(cap, dtype=np.uint8)
    n = C.c_size_t(0)
    rc = lib().ref_encode(_planes_ptr(planes), len(planes), w, h, prec, sgnd, C.byref(p),
                          C.c_void_p(out.ctypes.data), C.c_size_t(cap), C.byref(n))
    if rc != 0:
        raise RuntimeError("ref_encode failed rc=%d" % rc)
    return out[:n.value].tobytes()


def probe(data):
    info = (C.c_int * 5)()
    rc = lib().ref_decode(data, C.c_size_t(len(data)), 0, 1, None, 0, info)
    if rc != 0:
        raise RuntimeError("ref probe failed rc=%d" % rc)
    return dict(w=info[0], h=info[1], ncomp=info[2], prec=info[3], sgnd=info[4])


def decode(data, reduce=0, threads=1):
    """bytes -> list of (h,w) int32 arrays"""
    inf = probe(data)
    w = -(-inf["w"] // (1 << reduce))
    h = -(-inf["h"] // (1 << reduce))
    planes = [np.zeros((h, w), dtype=np.int32) for _ in range(inf["ncomp"])]
    info = (C.c_int * 5)()
    rc = lib().ref_decode(data, C.c_size_t(len(data)), reduce, threads, _planes_ptr(planes),
                          len(planes), info)
    if rc != 0:
        raise RuntimeError("ref_decode failed rc=%d" % rc)
    assert (info[0], info[1]) == (w, h), (info[0], info[1], w, h)
    return planes


def t1_encode_cblk(coeff, orient, compno=0, level=0, qmfbid=1, stepsize=1.0, use_mct_norms=0,
                   numcomps=1):
    coeff = np.ascontiguousarray(coeff, dtype=np.int32)
    h, w = coeff.shape
    out = np.empty(w * h * 4 + 64, dtype=np.uint8)
    numbps = C.c_int(0)
    npasses = C.c_int(0)
    rates = (C.c_int * 128)()
    dist = (C.c_double * 128)()
    rc = lib().ref_t1_encode_cblk(C.c_void_p(coeff.ctypes.data), w, h, w, orient, compno, level,
                                  qmfbid, C.c_double(stepsize), use_mct_norms, numcomps,
                                  C.c_void_p(out.ctypes.data), out.size, C.byref(numbps),
                                  C.byref(npasses), rates, dist)
    if rc != 0:
        raise RuntimeError("ref_t1_encode_cblk rc=%d" % rc)
    n = npasses.value
    r = [rates[i] for i in range(n)]
    total = r[-1] if n else 0
    return dict(data=out[:total].tobytes(), numbps=numbps.value, npasses=n, rates=r,
                dist=[dist[i] for i in range(n)])


def dwt_encode(plane, numres, real=0, x0=0, y0=0):
    a = np.ascontiguousarray(plane, dtype=np.int32).copy()
    h, w = a.shape
    rc = lib().ref_dwt_encode(C.c_void_p(a.ctypes.data), x0, y0, x0 + w, y0 + h, numres, real)
    assert rc == 0
    return a


def dwt_decode(plane, numres, real=0, x0=0, y0=0, threads=1):
    a = np.ascontiguousarray(plane).copy()
    assert a.dtype.itemsize == 4
    h, w = a.shape
    rc = lib().ref_dwt_decode(C.c_void_p(a.ctypes.data), x0, y0, x0 + w, y0 + h, numres, real, threads)
    assert rc == 0
    return a


def mct(c0, c1, c2, inverse=False, real=False):
    if real and inverse:
        a = [np.ascontiguousarray(c, dtype=np.float32).copy() for c in (c0, c1, c2)]
    else:
        a = [np.ascontiguousarray(c, dtype=np.int32).copy() for c in (c0, c1, c2)]
    fn = {(False, False): "ref_mct_encode", (True, False): "ref_mct_decode",
          (False, True): "ref_mct_encode_real", (True, True): "ref_mct_decode_real"}[(inverse, real)]
    getattr(lib(), fn)(C.c_void_p(a[0].ctypes.data), C.c_void_p(a[1].ctypes.data),
                       C.c_void_p(a[2].ctypes.data), C.c_uint64(a[0].size))
    return a


def set_extra(mode=0, roi_comp=-1, roi_shift=0, csty=0, sub=(1, 1)):
    """switches for the next encode() calls: code-block style bits (1 LAZY 2 RESET 4 TERMALL 8 VSC 16 PTERM 32 SEGSYM),
    region of interest, SOP (csty 2) / EPH (csty 4) markers, sub-sampling of components 1.."""
    lib().ref_set_extra(mode, roi_comp, roi_shift, csty, sub[0], sub[1])


def t1_decode_cblk(data, w, h, orient, numbps, segs, cblksty=0, roishift=0):
    """segs: list of (length, passes) -> (h, w) int32 in the decoder's representation (one extra fractional bit)"""
    buf = np.frombuffer(bytes(data) + b"\0" * 8, dtype=np.uint8).copy()
    sl = (C.c_int * len(segs))(*[s[0] for s in segs])
    sp = (C.c_int * len(segs))(*[s[1] for s in segs])
    out = np.zeros((h, w), dtype=np.int32)
    rc = lib().ref_t1_decode_cblk(C.c_void_p(buf.ctypes.data), sl, sp, len(segs), w, h, orient, numbps, roishift, cblksty,
                                  C.c_void_p(out.ctypes.data))
    if rc != 0:
        raise RuntimeError("ref_t1_decode_cblk rc=%d" % rc)
    return out


def t1_encode_cblk_generic(coeff, orient, cblksty):
    """the reference's generic encoder (used when cblksty != 0) -> bytes, numbps, per pass (rate, len, term)"""
    coeff = np.ascontiguousarray(coeff, dtype=np.int32)
    h, w = coeff.shape
    out = np.empty(w * h * 8 + 64, dtype=np.uint8)
    numbps = C.c_int(0); npasses = C.c_int(0)
    rates = (C.c_int * 128)(); lens = (C.c_int * 128)(); terms = (C.c_int * 128)()
    rc = lib().ref_t1_encode_cblk_generic(C.c_void_p(coeff.ctypes.data), w, h, w, orient, cblksty, C.c_void_p(out.ctypes.data), out.size,
                                          C.byref(numbps), C.byref(npasses), rates, lens, terms)
    if rc != 0:
        raise RuntimeError("ref_t1_encode_cblk_generic rc=%d" % rc)
    n = npasses.value
    total = rates[n - 1] if n else 0
    return dict(data=out[:total].tobytes(), numbps=numbps.value, npasses=n, rates=[rates[i] for i in range(n)],
                lens=[lens[i] for i in range(n)], terms=[terms[i] for i in range(n)])


def decode_components(data, reduce=0, threads=1):
    """like decode(), for streams whose components 1.. are sub-sampled: -> list of arrays, each at its component's own size"""
    inf = probe(data)
    dx, dy = C.c_int(1), C.c_int(1)
    lib().ref_last_subsampling(C.byref(dx), C.byref(dy))
    w = -(-inf["w"] // (1 << reduce)); h = -(-inf["h"] // (1 << reduce))
    planes = [np.zeros((h, w), dtype=np.int32) for _ in range(inf["ncomp"])]
    info = (C.c_int * 5)()
    rc = lib().ref_decode(data, C.c_size_t(len(data)), reduce, threads, _planes_ptr(planes), len(planes), info)
    if rc != 0:
        raise RuntimeError("ref_decode failed rc=%d" % rc)
    out = [planes[0]]
    for c in range(1, inf["ncomp"]):
        cw = -(-(-(-inf["w"] // dx.value)) // (1 << reduce))         # ceildivpow2(ceildiv(W, dx), reduce)
        chh = -(-(-(-inf["h"] // dy.value)) // (1 << reduce))
        out.append(planes[c].ravel()[:cw * chh].reshape(chh, cw).copy())
    return out


def encode_with_plugin(plugin_path, infile, irreversible=0, numresolution=6, cblk=(64, 64), mct=0, prog=0, numlayers=1, disto_alloc=1,
                       rsiz=0, max_comp_size=0, max_cs_size=0, rates=()):
    """the reference's accelerator path: opj_plugin_init + opj_plugin_encode + opj_encode_with_plugin (openjpeg.cpp:1077-1160)"""
    p = RefParams(irreversible, numresolution, cblk[0], cblk[1], mct, prog, 0, 0, numlayers, disto_alloc, rsiz, max_comp_size, max_cs_size, 1)
    for i, r in enumerate(rates):
        p.rates[i] = r
    cap = 64 << 20
    out = np.empty(cap, dtype=np.uint8)
    n = C.c_size_t(0)
    rc = lib().ref_encode_with_plugin(plugin_path.encode(), infile.encode(), C.byref(p), C.c_void_p(out.ctypes.data), C.c_size_t(cap), C.byref(n))
    if rc != 0:
        raise RuntimeError("ref_encode_with_plugin failed rc=%d" % rc)
    return out[:n.value].tobytes()
