"""Test-side helpers over the C ABI (ctypes): stage calls and whole-frame encode."""
import ctypes as C
import numpy as np
from j2k_b200 import _lib as L


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def stage_mct(c0, c1, c2, irreversible=0, inverse=0):
    a = [np.ascontiguousarray(c).copy() for c in (c0, c1, c2)]
    assert all(x.dtype.itemsize == 4 for x in a)
    rc = L.lib().j2kcuda_stage_mct(_vp(a[0]), _vp(a[1]), _vp(a[2]), C.c_size_t(a[0].size), irreversible, inverse)
    assert rc == 0, L.last_error()
    return a


def stage_dwt(plane, numres, irreversible=0, inverse=0, x0=0, y0=0):
    a = np.ascontiguousarray(plane).copy()
    assert a.dtype.itemsize == 4
    h, w = a.shape
    rc = L.lib().j2kcuda_stage_dwt(_vp(a), x0, y0, x0 + w, y0 + h, numres, irreversible, inverse)
    assert rc == 0, L.last_error()
    return a


def stage_t1_encode(plane, blocks, irreversible=0):
    """blocks: list of (x, y, w, h, orient, bandconst) -> list of dict per block"""
    plane = np.ascontiguousarray(plane, dtype=np.int32)
    h, w = plane.shape
    blk = np.array(blocks, dtype=np.int32).reshape(-1, 6)
    n = len(blk)
    numbps = np.zeros(n, np.int32); npasses = np.zeros(n, np.int32)
    rates = np.zeros((n, 128), np.int32); nms = np.zeros((n, 128), np.int32)
    cap = int(sum(int(b[2]) * int(b[3]) * 8 + 64 for b in blk))
    data = np.zeros(cap, np.uint8)
    offs = np.zeros(n + 1, np.uint64)
    rc = L.lib().j2kcuda_stage_t1_encode(_vp(plane), w, h, _vp(blk), n, irreversible, _vp(numbps), _vp(npasses),
                                         _vp(rates), _vp(nms), _vp(data), C.c_size_t(cap), _vp(offs))
    assert rc == 0, L.last_error()
    out = []
    for i in range(n):
        k = int(npasses[i])
        out.append(dict(numbps=int(numbps[i]), npasses=k, rates=[int(v) for v in rates[i, :k]],
                        nmsedec=[int(v) for v in nms[i, :k]], data=data[int(offs[i]):int(offs[i + 1])].tobytes()))
    return out


def planes_of(img):
    """(h, w, c) interleaved array -> ctypes array of j2kcuda_plane aliasing it (the AE layout)"""
    h, w, nc = img.shape
    S = img.dtype.itemsize
    arr = (L.Plane * nc)()
    for c in range(nc):
        arr[c].buf = img.ctypes.data + c * S
        arr[c].colbytes = nc * S
        arr[c].rowbytes = img.strides[0]
        arr[c].width = w; arr[c].height = h
        arr[c].sample_type = {1: L.UCHAR, 2: L.USHORT, 4: L.UINT}[S]
        arr[c].depth = 0; arr[c].sgnd = 0
    return arr


def encode(img, prec, **kw):
    h, w, nc = img.shape
    p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, **kw)
    pl = planes_of(img)
    out = C.POINTER(C.c_uint8)()
    n = C.c_size_t(0)
    rc = L.lib().j2kcuda_encode(C.byref(p), pl, nc, C.byref(out), C.byref(n))
    assert rc == 0, L.last_error()
    data = C.string_at(out, n.value)
    L.lib().j2kcuda_free(out)
    return data


def probe(data):
    info = L.Info()
    rc = L.lib().j2kcuda_probe(data, C.c_size_t(len(data)), C.byref(info))
    assert rc == 0, L.last_error()
    return info


def t2_parse(data):
    """-> list of (numbps, npasses, bytes) per code-block in Tier-2 order"""
    info = C.POINTER(C.c_int32)(); nb = C.c_int(0); seg = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
    rc = L.lib().j2kcuda_t2_parse(data, C.c_size_t(len(data)), 0, C.byref(info), C.byref(nb), C.byref(seg), C.byref(n))
    assert rc == 0, L.last_error()
    raw = C.string_at(seg, n.value)
    out = []
    for i in range(nb.value):
        numbps, npasses, ln, off = info[4 * i], info[4 * i + 1], info[4 * i + 2], info[4 * i + 3]
        out.append((numbps, npasses, raw[off:off + ln]))
    L.lib().j2kcuda_free(info); L.lib().j2kcuda_free(seg)
    return out


def stage_t1_decode(blocks):
    """blocks: list of (w, h, orient, numbps, npasses, bytes) -> list of (h, w) int32 arrays (reference
    decoder representation: one extra fractional bit)"""
    data = b"".join(b[5] for b in blocks)
    blk = np.array([[b[0], b[1], b[2], b[3], b[4], len(b[5])] for b in blocks], dtype=np.int32)
    total = int(sum(b[0] * b[1] for b in blocks))
    out = np.zeros(max(total, 1), np.int32)
    buf = np.frombuffer(data + b"\0", dtype=np.uint8).copy()
    rc = L.lib().j2kcuda_stage_t1_decode(_vp(buf), _vp(blk), len(blocks), _vp(out))
    assert rc == 0, L.last_error()
    res, pos = [], 0
    for b in blocks:
        res.append(out[pos:pos + b[0] * b[1]].reshape(b[1], b[0]).copy())
        pos += b[0] * b[1]
    return res


def decode(data, reduce=0, dtype=None):
    """codestream -> (h, w, ncomp) array through j2kcuda_decode (host destination, interleaved)"""
    info = probe(data)
    w = -(-info.width // (1 << reduce)); h = -(-info.height // (1 << reduce)); nc = info.ncomp
    if dtype is None:
        dtype = np.uint8 if max(info.prec[c] for c in range(nc)) <= 8 else np.uint16
    img = np.zeros((h, w, nc), dtype=dtype)
    pl = planes_of(img)
    rc = L.lib().j2kcuda_decode(data, C.c_size_t(len(data)), reduce, pl, nc, None, None)
    assert rc == 0, L.last_error()
    return img
