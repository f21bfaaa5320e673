"""Test helper: whole-frame encode built from the ORACLE's stage functions (oracle/liboracle.so) plus
the product's host Tier-2 (j2kcuda_t2_assemble).  Lets the CPU-only suite check the Tier-2 writer
byte for byte against the compiled reference without a GPU.  Test infrastructure only."""
import ctypes as C
import numpy as np
import orc
from j2k_b200 import _lib as L


def block_list(params):
    """geometry of every code-block, in Tier-2 order, via the library's own table"""
    n = L.lib().j2kcuda_t2_block_count(C.byref(params))
    assert n >= 0
    geo = np.zeros((n, 12), dtype=np.int32)
    got = L.lib().j2kcuda_t2_block_geometry(C.byref(params), geo.ctypes.data_as(C.c_void_p), n)
    assert got == n, (got, n)
    return geo


def oracle_tile_stages(planes, prec, sgnd, params, x0=0, y0=0):
    """planes: list of int32 (h,w) unsigned samples of ONE tile whose top-left sample is (x0, y0) of the image (the parity
    of the origin at every level decides which samples are low-pass, dwt.cpp:420-506).  Returns DWT coefficient planes."""
    irreversible = int(params.irreversible)
    numres = int(params.numresolution) or 6
    shift = 0 if sgnd else 1 << (prec - 1)
    p = [np.ascontiguousarray(x, dtype=np.int32) - shift for x in planes]
    if irreversible:
        p = [x * 2048 for x in p]          # tcd.cpp:1659-1696: 11 fractional bits
    if params.mct and len(p) >= 3:
        p[:3] = orc.mct(p[0], p[1], p[2], real=bool(irreversible))
    return [orc.dwt(x, numres, real=bool(irreversible), x0=x0, y0=y0) for x in p]


def encode(planes, prec, params, sgnd=0):
    """whole frame -> codestream bytes, using oracle stages + product Tier-2"""
    lib = L.lib()
    geo = block_list(params)
    h, w = planes[0].shape
    tw = int(params.tile_w) or w
    th = int(params.tile_h) or h
    ntx = -(-w // tw)
    irreversible = int(params.irreversible)
    coeffs = {}
    results = (L.BlockResult * len(geo))()
    keep = []
    for i, g in enumerate(geo):
        tile, comp, px, py, bw, bh, orient, level, numbps_band, bandconst, stepsize_bits, _ = [int(v) for v in g]
        if tile not in coeffs:
            tx, ty = tile % ntx, tile // ntx
            sub = [pl[ty * th:(ty + 1) * th, tx * tw:(tx + 1) * tw] for pl in planes]
            coeffs[tile] = oracle_tile_stages(sub, prec, sgnd, params, x0=tx * tw, y0=ty * th)
        blk = coeffs[tile][comp][py:py + bh, px:px + bw]
        if bw == 0 or bh == 0:
            r = dict(data=b"", numbps=0, npasses=0, rates=[], nmsedec=[])
        else:
            q, mx = orc.quantize(blk, qmfbid=0 if irreversible else 1, bandconst=bandconst)
            sty = int(params.cblksty)
            r = orc.t1_encode_sty(q, orient, sty, mx) if sty else orc.t1_encode(q, orient, mx)
        stepsize = np.array([stepsize_bits], dtype=np.int32).view(np.float32)[0]
        mctn = 1.0
        if params.mct and comp < 3 and int(params.ncomp) >= 3:
            mctn = orc.lib().orc_mct_norm(comp, irreversible)
        dn = orc.lib().orc_dwt_norm(level, orient, irreversible)
        dist = orc.distortions(r["nmsedec"], r["numbps"], mctn, dn, float(stepsize))
        data = np.frombuffer(r["data"] + b"\0", dtype=np.uint8).copy()
        rates = np.array(r["rates"] + [0], dtype=np.int32)
        dd = np.array(dist + [0.0], dtype=np.float64)
        keep.append((data, rates, dd))
        results[i].data = data.ctypes.data_as(C.POINTER(C.c_uint8))
        results[i].numbps = r["numbps"]
        results[i].npasses = r["npasses"]
        results[i].rates = rates.ctypes.data_as(C.POINTER(C.c_int32))
        results[i].dist = dd.ctypes.data_as(C.POINTER(C.c_double))
    out = C.POINTER(C.c_uint8)()
    n = C.c_size_t(0)
    rc = lib.j2kcuda_t2_assemble(C.byref(params), results, len(geo), C.byref(out), C.byref(n))
    assert rc == 0, L.last_error()
    data = C.string_at(out, n.value)
    lib.j2kcuda_free(out)
    return data
