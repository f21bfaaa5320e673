"""Edge cases of the codec path, against the compiled reference: degenerate and ragged geometries (1x1, single row /
column, smaller than 2^levels, widths that are not a multiple of the 16-byte vector), empty content (all-zero planes:
every code-block absent from the packets), saturated content, non-square code-blocks, two components, odd precisions,
ragged tiles; and the error behaviour of the C ABI (bad parameters, truncated and corrupt streams: a status and a
message, never a crash).  CPU: host Tier-2 + oracle stages.  GPU: the CUDA path through the C ABI."""
import ctypes as C
import numpy as np
import pytest
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame
import hostpipe
import refgrok

needs_ref = pytest.mark.skipif(not refgrok.available(), reason="oracle/_ref not built")

GEOMETRIES = [  # h, w, ncomp, prec, numres, cblk_w, cblk_h, tile
    (1, 1, 1, 8, 1, 64, 64, 0), (1, 1, 3, 8, 6, 64, 64, 0), (1, 37, 1, 8, 6, 64, 64, 0), (41, 1, 1, 16, 6, 64, 64, 0),
    (2, 2, 3, 8, 3, 64, 64, 0), (3, 5, 3, 12, 6, 64, 64, 0), (5, 7, 4, 16, 6, 32, 32, 0), (17, 130, 3, 8, 5, 64, 16, 0),
    (70, 33, 2, 10, 4, 16, 64, 0), (65, 129, 3, 8, 2, 4, 4, 0), (50, 75, 3, 9, 4, 64, 64, 32), (33, 100, 1, 1, 3, 64, 64, 0),
]


def content(kind, h, w, nc, prec):
    if kind == "zero":
        return np.zeros((h, w, nc), np.uint16 if prec > 8 else np.uint8)
    if kind == "max":
        return np.full((h, w, nc), (1 << prec) - 1, np.uint16 if prec > 8 else np.uint8)
    if kind == "mid":                       # exactly the DC level: every coefficient zero after the level shift
        return np.full((h, w, nc), 1 << (prec - 1), np.uint16 if prec > 8 else np.uint8)
    return make_frame(h, w, nc, prec, kind, seed=11)


def ref_encode(img, prec, numres, cw, ch, tile, mct):
    """-> bytes, or None when the reference refuses the parameters (e.g. more resolutions than the tile has samples for)"""
    planes = [img[..., c].astype(np.int32) for c in range(img.shape[2])]
    try:
        return refgrok.encode(planes, prec, mct=mct, numresolution=numres, cblk=(cw, ch), tile=(tile, tile))
    except RuntimeError:
        return None


@needs_ref
@pytest.mark.parametrize("h,w,nc,prec,numres,cw,ch,tile", GEOMETRIES)
@pytest.mark.parametrize("kind", ["noise", "zero"])
def test_host_tier2_on_degenerate_geometries(h, w, nc, prec, numres, cw, ch, tile, kind):
    img = content(kind, h, w, nc, prec)
    mct = 1 if nc >= 3 else 0
    p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, mct=mct, numresolution=numres, cblk_w=cw, cblk_h=ch, tile_w=tile, tile_h=tile)
    ref = ref_encode(img, prec, numres, cw, ch, tile, mct)
    if ref is None:
        if numres > 1 and min(tile or w, tile or h) < (1 << (numres - 1)):
            # what the reference refuses, the library refuses (with the reference's message, j2k.cpp:6664-6672)
            assert L.lib().j2kcuda_t2_block_count(C.byref(p)) < 0
            assert "resolutions" in L.last_error()
            return
        pytest.skip("the reference's tile buffer (0.1625 x raw size, j2k.cpp:4451-4462) overflows on this input")
    ours = hostpipe.encode([img[..., c].astype(np.int32) for c in range(nc)], prec, p)
    assert ours == ref


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("h,w,nc,prec,numres,cw,ch,tile", GEOMETRIES)
@pytest.mark.parametrize("kind", ["noise", "zero", "max", "mid"])
def test_cuda_path_on_degenerate_geometries(h, w, nc, prec, numres, cw, ch, tile, kind):
    import j2k
    img = content(kind, h, w, nc, prec)
    mct = 1 if nc >= 3 else 0
    ref = ref_encode(img, prec, numres, cw, ch, tile, mct)
    if ref is None and numres > 1 and min(tile or w, tile or h) < (1 << (numres - 1)):
        p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, mct=mct, numresolution=numres, cblk_w=cw, cblk_h=ch, tile_w=tile, tile_h=tile)
        L.lib().j2kcuda_encoder_create.restype = C.c_void_p
        assert not L.lib().j2kcuda_encoder_create(C.byref(p))
        return
    ours = j2k.encode(img, prec, mct=mct, numresolution=numres, cblk_w=cw, cblk_h=ch, tile_w=tile, tile_h=tile)
    if ref is not None:
        assert ours == ref
    else:                                   # the reference's encoder overflows its own tile buffer here; its decoder is the arbiter
        assert all(np.array_equal(a, img[..., c]) for c, a in enumerate(refgrok.decode(ours)))
    assert np.array_equal(j2k.decode(ours), img)
    for reduce in (1, 2):
        if reduce < numres:
            want = refgrok.decode(ours, reduce=reduce)
            got = j2k.decode(ours, reduce=reduce)
            assert all(np.array_equal(got[..., c], want[c]) for c in range(nc))


@pytest.mark.gpu
def test_code_blocks_wider_or_taller_than_64_are_refused_not_miscoded():
    """the reference allows up to 1024 x 4; the CUDA Tier-1 keeps a block's rows as 64-bit masks"""
    lib = L.lib()
    lib.j2kcuda_encoder_create.restype = C.c_void_p
    for cw, ch in ((128, 32), (32, 128), (1024, 4)):
        p = L.default_params(width=256, height=256, ncomp=1, prec=8, cblk_w=cw, cblk_h=ch)
        assert not lib.j2kcuda_encoder_create(C.byref(p))
        assert "64x64" in L.last_error()


NARROW = [(40, 49, 1, 6, 48), (40, 52, 1, 6, 48), (40, 100, 3, 6, 48), (37, 50, 1, 5, 48), (49, 40, 1, 6, 48), (70, 130, 3, 6, 64)]


@needs_ref
@pytest.mark.parametrize("h,w,nc,numres,tile", NARROW)
def test_tiles_narrower_than_the_transform_depth(h, w, nc, numres, tile):
    """Bands of zero width: the reference's encoder writes bits for phantom code-blocks its own decoder then skips, so it
    cannot read its own stream back (shown here); the library writes what OpenJPEG 2.5.4 writes, and every decoder reads it."""
    import io
    Image = pytest.importorskip("PIL.Image")
    img = make_frame(h, w, nc, 8, "noise", seed=11).astype(np.uint8)
    planes = [img[..., c].astype(np.int32) for c in range(nc)]
    mct = 1 if nc == 3 else 0
    ref = refgrok.encode(planes, 8, mct=mct, numresolution=numres, tile=(tile, tile))
    try:
        ref_ok = all(np.array_equal(a, b) for a, b in zip(refgrok.decode(ref), planes))
    except RuntimeError:
        ref_ok = False
    assert not ref_ok, "the reference decodes its own stream here: the departure from byte identity is no longer needed"
    p = L.default_params(width=w, height=h, ncomp=nc, prec=8, mct=mct, numresolution=numres, tile_w=tile, tile_h=tile)
    ours = hostpipe.encode(planes, 8, p)
    assert all(np.array_equal(a, b) for a, b in zip(refgrok.decode(ours), planes))            # grok's decoder reads ours
    back = np.array(Image.open(io.BytesIO(ours)))
    assert np.array_equal(back.reshape(h, w, nc), img)                                        # so does OpenJPEG 2.5.4
    buf = io.BytesIO()
    Image.fromarray(img.squeeze() if nc == 1 else img, "L" if nc == 1 else "RGB").save(
        buf, "JPEG2000", no_jp2=True, irreversible=False, num_resolutions=numres, codeblock_size=(64, 64), mct=mct,
        progression="LRCP", tile_size=(tile, tile))
    p.compat = L.COMPAT_OPENJPEG
    assert hostpipe.encode(planes, 8, p) == buf.getvalue()                                    # and it is what OpenJPEG writes


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("h,w,nc,numres,tile", NARROW)
def test_cuda_path_on_tiles_narrower_than_the_transform_depth(h, w, nc, numres, tile):
    import j2k
    for prec, kind in ((8, "noise"), (16, "smooth")):
        img = make_frame(h, w, nc, prec, kind, seed=11)
        mct = 1 if nc == 3 else 0
        ours = j2k.encode(img, prec, mct=mct, numresolution=numres, tile_w=tile, tile_h=tile)
        assert np.array_equal(j2k.decode(ours), img)
        assert all(np.array_equal(a, img[..., c]) for c, a in enumerate(refgrok.decode(ours)))


def test_bad_parameters_are_refused_with_a_message():
    lib = L.lib()
    lib.j2kcuda_encoder_create.restype = C.c_void_p
    for kw in (dict(width=0, height=4, ncomp=1, prec=8), dict(width=4, height=4, ncomp=0, prec=8), dict(width=4, height=4, ncomp=5, prec=8),
               dict(width=4, height=4, ncomp=1, prec=0), dict(width=4, height=4, ncomp=1, prec=17),
               dict(width=4, height=4, ncomp=1, prec=8, cblk_w=48), dict(width=4, height=4, ncomp=1, prec=8, cblk_w=1024, cblk_h=1024),
               dict(width=4, height=4, ncomp=1, prec=8, numresolution=40)):
        p = L.default_params(**kw)
        assert lib.j2kcuda_t2_block_count(C.byref(p)) < 0 or not lib.j2kcuda_encoder_create(C.byref(p)), kw
    assert lib.j2kcuda_probe(None, C.c_size_t(0), None) == L.ERR_PARAM
    info = L.Info()
    assert lib.j2kcuda_probe(b"", C.c_size_t(0), C.byref(info)) in (L.ERR_STREAM, L.ERR_PARAM)


@needs_ref
def test_truncated_and_corrupt_streams_are_rejected_by_the_parser():
    img = make_frame(40, 60, 3, 8, "smooth", seed=1)
    cs = ref_encode(img, 8, 4, 64, 64, 0, 1)
    lib = L.lib()
    info = C.POINTER(C.c_int32)(); nb = C.c_int(0); seg = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)

    def parse(data):
        rc = lib.j2kcuda_t2_parse(data, C.c_size_t(len(data)), 0, C.byref(info), C.byref(nb), C.byref(seg), C.byref(n))
        if rc == 0:
            lib.j2kcuda_free(info); lib.j2kcuda_free(seg)
        return rc

    assert parse(cs) == 0
    rng = np.random.default_rng(0)
    for cut in (2, 10, 40, 60, 100, len(cs) // 2, len(cs) - 3):
        rc = parse(cs[:cut])
        assert rc in (0, L.ERR_STREAM, L.ERR_UNSUPPORTED)            # the reference also accepts a stream cut inside the last packets
        if cut < 100:
            assert rc != 0 and L.last_error()
    for _ in range(200):                                             # random byte damage in the headers: a status, never a crash
        bad = bytearray(cs)
        for k in rng.integers(2, 120, size=3):
            bad[int(k)] = int(rng.integers(0, 256))
        assert parse(bytes(bad)) in (0, L.ERR_STREAM, L.ERR_UNSUPPORTED, L.ERR_PARAM)


# ---- crafted headers and packets: a status and a message, never a fault, an abort or an over-read ----
def _hdr(w=64, h=64, tdx=None, tdy=None, x0=0, y0=0, tx0=0, ty0=0, ncomp=1, layers=1, numres=1, csty=0, prec_bytes=b"", qcd=None):
    import struct
    tdx = w if tdx is None else tdx
    tdy = h if tdy is None else tdy
    siz = struct.pack(">HIIIIIIIIH", 0, w, h, x0, y0, tdx, tdy, tx0, ty0, ncomp) + b"\x07\x01\x01" * ncomp
    cod = struct.pack(">BBHB", csty, 0, layers, 0) + bytes([numres - 1, 4, 4, 0, 1]) + prec_bytes
    qcd = bytes([0x40]) + bytes([8 << 3] * (3 * numres - 2)) if qcd is None else qcd
    seg = lambda m, b: struct.pack(">HH", m, len(b) + 2) + b
    return b"\xff\x4f" + seg(0xff51, siz) + seg(0xff52, cod) + seg(0xff5c, qcd)


def _parse(data):
    lib = L.lib()
    info = C.POINTER(C.c_int32)(); nb = C.c_int(0); seg = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
    rc = lib.j2kcuda_t2_parse(data, C.c_size_t(len(data)), 0, C.byref(info), C.byref(nb), C.byref(seg), C.byref(n))
    if rc == 0:
        lib.j2kcuda_free(info); lib.j2kcuda_free(seg)
    return rc


def _tile(body):
    import struct
    return b"\xff\x90" + struct.pack(">HHIBB", 10, 0, 14 + len(body), 0, 1) + b"\xff\x93" + body + b"\xff\xd9"


def test_segment_length_of_32_one_bits_is_rejected():
    # packet header: non-empty, block included, 0 missing bit-planes, 1 pass, Lblock grown by 29 (3 + 29 = 32 bits),
    # then a length of 32 one-bits: read as int this was -1 and passed the bounds check (heap overflow, ADVICE high)
    bits = "1" + "1" + "1" + "0" + "1" * 29 + "0" + "1" * 32
    by = bytearray()
    acc, cnt, lim = 0, 0, 8
    for b in bits + "0" * 16:                          # bit-stuffed like bio.cpp: 7 bits after a 0xff byte
        acc = (acc << 1) | int(b); cnt += 1
        if cnt == lim:
            by.append(acc); lim = 7 if acc == 0xff else 8; acc = 0; cnt = 0
    data = _hdr() + _tile(bytes(by) + b"\x00" * 64)
    assert _parse(data) == L.ERR_STREAM and L.last_error()


def test_giant_and_degenerate_geometry_is_refused_before_any_allocation():
    for hdr in (_hdr(w=1 << 30, h=1 << 30),                                  # 2^60 samples
                _hdr(w=1 << 20, h=1 << 10, numres=1, csty=1, prec_bytes=b"\x00"),   # one-sample precincts over a megapixel
                _hdr(w=1024, h=1024, layers=65535, numres=6, csty=1, prec_bytes=b"\x22" * 6),   # 65535 layers x 87 000 precincts
                _hdr(w=64, h=64, x0=40, y0=40, tdx=16, tdy=16),               # first tile ends before the image starts
                _hdr(w=64, h=64, numres=3, csty=1, prec_bytes=b"\x88\x80\x88")):   # precinct exponent 0 above resolution 0 (j2k.cpp:8445)
        assert _parse(hdr + _tile(b"\x00" * 16)) in (L.ERR_STREAM, L.ERR_NOMEM, L.ERR_UNSUPPORTED), L.last_error()
        assert L.last_error()


def test_marker_segments_shorter_than_their_contents():
    import struct
    good = _hdr()
    assert _parse(good + _tile(b"\x80" + b"\x00" * 8)) in (0, L.ERR_STREAM)
    siz_end = 2 + 4 + 38 + 3
    for marker, L_ in ((0xff52, 3), (0xff52, 11), (0xff53, 4), (0xff5c, 3), (0xff5d, 3)):
        bad = good[:siz_end] + struct.pack(">HH", marker, L_) + b"\x00" * (L_ - 2)     # segment ends the buffer: the old reader ran past it
        assert _parse(bad) == L.ERR_STREAM and L.last_error()
