"""ctypes wrapper over oracle/liboracle.so (our C restatement).  Test infrastructure only."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_DIR = os.path.join(_HERE, "..", "oracle")
_PATH = os.path.join(_DIR, "liboracle.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            subprocess.check_call(["make", "-C", _DIR, "oracle"])
        _lib = C.CDLL(_PATH)
        _lib.orc_wmsedec.restype = C.c_double
        _lib.orc_wmsedec.argtypes = [C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
        _lib.orc_dwt_norm.restype = C.c_double
        _lib.orc_mct_norm.restype = C.c_double
        _lib.orc_quantize_block.restype = C.c_uint32
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data)


def mct(c0, c1, c2, inverse=False, real=False):
    dt = np.float32 if (real and inverse) else np.int32
    a = [np.ascontiguousarray(c, dtype=dt).copy() for c in (c0, c1, c2)]
    fn = {(False, False): "orc_rct_fwd", (True, False): "orc_rct_inv",
          (False, True): "orc_ict_fwd", (True, True): "orc_ict_inv"}[(inverse, real)]
    getattr(lib(), fn)(_p(a[0]), _p(a[1]), _p(a[2]), C.c_size_t(a[0].size))
    return a


def dwt(plane, numres, inverse=False, real=False, x0=0, y0=0):
    dt = np.float32 if (real and inverse) else np.int32
    a = np.ascontiguousarray(plane, dtype=dt).copy()
    h, w = a.shape
    fn = {(False, False): "orc_dwt53_fwd", (True, False): "orc_dwt53_inv",
          (False, True): "orc_dwt97_fwd", (True, True): "orc_dwt97_inv"}[(inverse, real)]
    getattr(lib(), fn)(_p(a), x0, y0, x0 + w, y0 + h, numres)
    return a


def quantize(coeff, qmfbid=1, bandconst=0):
    coeff = np.ascontiguousarray(coeff, dtype=np.int32)
    h, w = coeff.shape
    out = np.empty((h, w), dtype=np.uint32)
    mx = lib().orc_quantize_block(_p(coeff), w, w, h, qmfbid, bandconst, _p(out))
    return out, mx


def t1_encode(smag, orient, maxmag=None):
    smag = np.ascontiguousarray(smag, dtype=np.uint32)
    h, w = smag.shape
    if maxmag is None:
        maxmag = int((smag & 0x7FFFFFFF).max()) if smag.size else 0
    out = np.empty(w * h * 4 + 64, dtype=np.uint8)
    numbps = C.c_int(0)
    npasses = C.c_int(0)
    rates = (C.c_int * 128)()
    nms = (C.c_int * 128)()
    n = lib().orc_t1_encode_cblk(_p(smag), w, h, orient, C.c_uint32(maxmag), _p(out), out.size,
                                 C.byref(numbps), C.byref(npasses), rates, nms)
    assert n >= 0
    k = npasses.value
    return dict(data=out[:n].tobytes(), numbps=numbps.value, npasses=k,
                rates=[rates[i] for i in range(k)], nmsedec=[nms[i] for i in range(k)])


def t1_encode_sty(smag, orient, cblksty, maxmag=None, trace=False):
    """the generic encoder loop (RESET 2 / TERMALL 4): like t1_encode; with trace also the (ctx << 1 | d) symbols and per-pass counts"""
    smag = np.ascontiguousarray(smag, dtype=np.uint32)
    h, w = smag.shape
    if maxmag is None:
        maxmag = int((smag & 0x7FFFFFFF).max()) if smag.size else 0
    out = np.empty(w * h * 8 + 64, dtype=np.uint8)
    numbps = C.c_int(0)
    npasses = C.c_int(0)
    rates = (C.c_int * 128)()
    nms = (C.c_int * 128)()
    tr = np.zeros(w * h * 64 + 64, dtype=np.uint8) if trace else None
    pn = (C.c_uint32 * 128)()
    n = lib().orc_t1_encode_cblk_sty(_p(smag), w, h, orient, C.c_uint32(maxmag), cblksty, _p(out), out.size, C.byref(numbps),
                                     C.byref(npasses), rates, nms, _p(tr) if trace else None, tr.size if trace else 0, pn)
    assert n >= 0, n
    k = npasses.value
    r = dict(data=out[:n].tobytes(), numbps=numbps.value, npasses=k, rates=[rates[i] for i in range(k)], nmsedec=[nms[i] for i in range(k)])
    if trace:
        r["pass_nsym"] = [pn[i] for i in range(k)]
        r["symbols"] = tr[:sum(r["pass_nsym"])].copy()
    return r


def t1_decode(data, w, h, orient, numbps, npasses):
    out = np.zeros((h, w), dtype=np.int32)
    buf = np.frombuffer(data, dtype=np.uint8)
    buf = np.ascontiguousarray(buf)
    rc = lib().orc_t1_decode_cblk(_p(buf) if buf.size else None, len(data), w, h, orient, numbps,
                                  npasses, _p(out))
    assert rc == 0
    return out


def distortions(nmsedec, numbps, mct_norm, dwt_norm, stepsize):
    """cumulative distortion per pass, t1_opt.cpp:711-722"""
    out = []
    cum = 0.0
    bpno = numbps - 1
    passtype = 2
    for nm in nmsedec:
        cum += lib().orc_wmsedec(nm, mct_norm, dwt_norm, stepsize, bpno)
        out.append(cum)
        passtype += 1
        if passtype == 3:
            passtype = 0
            bpno -= 1
    return out
