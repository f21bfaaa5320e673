"""JP2 file wrapper (host only, no GPU): j2kcuda_wrap_jp2 is byte-identical to what the compiled reference
(ext/grok jp2.cpp, via oracle/_ref) writes around the same codestream; j2kcuda_probe_jp2 reads the header boxes of
files written by the reference, by OpenJPEG 2.5 (Pillow) and of hand-built files with ICC profile, palette,
component mapping, per-component depths and resolution boxes."""
import ctypes as C
import io
import struct
import numpy as np
import pytest
from j2k_b200 import _lib as L
import refgrok

needs_ref = pytest.mark.skipif(not refgrok.available(), reason="oracle/_ref not built")


def probe(data):
    info = L.Jp2Info()
    rc = L.lib().j2kcuda_probe_jp2(data, C.c_size_t(len(data)), C.byref(info))
    return rc, info


def wrap(cs, w, h, nc, prec, sgnd=0, enumcs=16, icc=None, alpha=-1):
    p = L.Jp2Params()
    p.width, p.height, p.ncomp, p.enumcs, p.alpha_comp = w, h, nc, enumcs, alpha
    precs = prec if isinstance(prec, (list, tuple)) else [prec] * nc
    for i in range(nc):
        p.prec[i] = precs[i]; p.sgnd[i] = sgnd
    if icc:
        p.icc = icc; p.icc_len = len(icc)
    out = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
    rc = L.lib().j2kcuda_wrap_jp2(cs, C.c_size_t(len(cs)), C.byref(p), C.byref(out), C.byref(n))
    assert rc == 0, L.last_error()
    data = C.string_at(out, n.value)
    L.lib().j2kcuda_free(out)
    return data


def box(t, body):
    return struct.pack(">I4s", 8 + len(body), t) + body


def ref_pair(nc, prec, alpha, h=40, w=50):
    rng = np.random.default_rng(nc * 100 + prec)
    pl = [rng.integers(0, 1 << prec, size=(h, w)).astype(np.int32) for _ in range(nc)]
    lib = refgrok.lib()
    lib.ref_set_jp2(1, alpha)
    try:
        jp2 = refgrok.encode(pl, prec, mct=1 if nc >= 3 else 0)
    finally:
        lib.ref_set_jp2(0, -1)
    cs = refgrok.encode(pl, prec, mct=1 if nc >= 3 else 0)
    return jp2, cs, w, h


@needs_ref
@pytest.mark.parametrize("nc,prec,alpha", [(3, 8, -1), (4, 8, 3), (4, 16, 3), (1, 8, -1), (2, 12, 1), (4, 8, -1), (4, 8, 1)])
def test_wrap_is_byte_identical_to_the_reference(nc, prec, alpha):
    jp2, cs, w, h = ref_pair(nc, prec, alpha)
    assert jp2.endswith(cs)
    ours = wrap(cs, w, h, nc, prec, enumcs=16 if nc >= 3 else 17, alpha=alpha)
    assert ours == jp2


@needs_ref
def test_probe_reads_the_reference_file():
    jp2, cs, w, h = ref_pair(4, 16, 3)
    rc, i = probe(jp2)
    assert rc == 0, L.last_error()
    assert (i.is_jp2, i.brand, i.width, i.height, i.ncomp, i.bpc) == (1, 0x6a703220, w, h, 4, 15)
    assert (i.colr_method, i.enumcs, i.cdef_n) == (1, 16, 4)
    assert [(i.cdef_cn[k], i.cdef_typ[k], i.cdef_asoc[k]) for k in range(4)] == [(0, 0, 1), (1, 0, 2), (2, 0, 3), (3, 1, 0)]
    assert jp2[i.codestream_offset:i.codestream_offset + i.codestream_length] == cs
    assert list(i.bpcc[:4]) == [15] * 4


def test_probe_raw_codestream_and_garbage():
    rc, i = probe(b"\xff\x4f\xff\x51" + b"\0" * 40)
    assert rc == 0 and i.is_jp2 == 0 and i.codestream_length == 44
    rc, _ = probe(b"not a jpeg 2000 file at all")
    assert rc == L.ERR_STREAM
    rc, _ = probe(box(b"jP  ", b"\x0d\x0a\x87\x0a") + box(b"ftyp", b"jp2 \0\0\0\0jp2 "))
    assert rc == L.ERR_STREAM and "codestream box" in L.last_error()


def test_probe_reads_openjpeg_2_5_files():
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(1)
    for mode, nc in (("RGBA", 4), ("RGB", 3), ("L", 1), ("LA", 2)):
        a = rng.integers(0, 255, size=(33, 47, nc)).astype(np.uint8)
        buf = io.BytesIO()
        Image.fromarray(a.squeeze() if nc == 1 else a, mode).save(buf, "JPEG2000", no_jp2=False)
        data = buf.getvalue()
        rc, i = probe(data)
        assert rc == 0, L.last_error()
        assert (i.width, i.height, i.ncomp, i.bpc, i.colr_method) == (47, 33, nc, 7, 1)
        assert i.enumcs == (16 if nc >= 3 else 17)
        assert data[i.codestream_offset:i.codestream_offset + 2] == b"\xff\x4f"
        if nc in (2, 4):
            assert i.cdef_n == nc and i.cdef_typ[nc - 1] == 1 and i.cdef_asoc[nc - 1] == 0


def test_probe_icc_palette_cmap_bpcc_res():
    icc = bytes(range(200)) * 2
    pal = bytes([r for e in range(5) for r in (e, 2 * e, 255 - e)])
    ihdr = struct.pack(">IIHBBBB", 20, 30, 1, 255, 7, 1, 0)
    resc = box(b"resc", struct.pack(">HHHHbb", 300, 1, 150, 2, 2, 1))     # 300e2 vertical, 75e1 horizontal
    jp2h = (box(b"ihdr", ihdr) + box(b"bpcc", bytes([7])) + box(b"colr", bytes([2, 0, 0]) + icc) +
            box(b"colr", bytes([1, 0, 0]) + struct.pack(">I", 17)) +      # a second colr box is ignored (jp2.cpp:1660-1667)
            box(b"pclr", struct.pack(">HB", 5, 3) + bytes([7, 7, 7]) + pal) +
            box(b"cmap", b"".join(struct.pack(">HBB", 0, 1, c) for c in range(3))) + box(b"res ", resc))
    cs = b"\xff\x4f\xff\x51" + b"\0" * 10
    data = (box(b"jP  ", b"\x0d\x0a\x87\x0a") + box(b"ftyp", b"jpx \0\0\0\0jp2 jpx ") + box(b"xml ", b"<a/>") +
            box(b"jp2h", jp2h) + struct.pack(">I4s", 0, b"jp2c") + cs)     # LBox 0: box runs to the end of the file
    rc, i = probe(data)
    assert rc == 0, L.last_error()
    assert (i.brand, i.ncomp, i.bpc, i.bpcc[0], i.unk_c) == (0x6a707820, 1, 255, 7, 1)
    assert (i.colr_method, i.icc_length) == (2, len(icc))
    assert data[i.icc_offset:i.icc_offset + i.icc_length] == icc
    assert (i.pclr_entries, i.pclr_channels, list(i.pclr_depth[:3])) == (5, 3, [7, 7, 7])
    assert [(i.cmap_cmp[k], i.cmap_mtyp[k], i.cmap_pcol[k]) for k in range(3)] == [(0, 1, 0), (0, 1, 1), (0, 1, 2)]
    assert i.has_capture_res == 1 and i.has_display_res == 0
    assert abs(i.capture_res[0] - 30000.0) < 1e-9 and abs(i.capture_res[1] - 750.0) < 1e-9
    assert data[i.codestream_offset:] == cs and i.codestream_length == len(cs)
    ent = (C.c_uint32 * 15)()
    n = L.lib().j2kcuda_jp2_palette(data, C.c_size_t(len(data)), ent, C.c_size_t(15))
    assert n == 15 and list(ent) == list(pal)
    # the round trip through our own writer keeps the ICC profile
    w = wrap(cs, 30, 20, 3, 8, icc=icc)
    rc, j = probe(w)
    assert rc == 0 and j.colr_method == 2 and w[j.icc_offset:j.icc_offset + j.icc_length] == icc and j.cdef_n == 0


def test_wrap_writes_bpcc_for_mixed_depths():
    cs = b"\xff\x4f\xff\x51" + b"\0" * 6
    w = wrap(cs, 8, 8, 3, [8, 10, 8])
    rc, i = probe(w)
    assert rc == 0 and i.bpc == 255 and list(i.bpcc[:3]) == [7, 9, 7]


def test_truncated_and_inconsistent_boxes_are_rejected():
    ihdr = struct.pack(">IIHBBBB", 20, 30, 3, 255, 7, 0, 0)
    head = box(b"jP  ", b"\x0d\x0a\x87\x0a") + box(b"ftyp", b"jp2 \0\0\0\0jp2 ")
    rc, _ = probe(head + box(b"jp2h", box(b"ihdr", ihdr)) + box(b"jp2c", b"\xff\x4f"))
    assert rc == L.ERR_STREAM and "bpcc" in L.last_error()
    rc, _ = probe(head + box(b"jp2h", box(b"ihdr", ihdr)[:-3]) + box(b"jp2c", b"\xff\x4f"))
    assert rc == L.ERR_STREAM
    rc, _ = probe(head + box(b"jp2c", b"\xff\x4f"))
    assert rc == L.ERR_STREAM and "before jp2h" in L.last_error()
    ok = struct.pack(">IIHBBBB", 20, 30, 1, 7, 7, 0, 0)
    rc, _ = probe(head + box(b"jp2h", box(b"ihdr", ok) + box(b"cmap", struct.pack(">HBB", 0, 1, 0))) + box(b"jp2c", b"\xff\x4f"))
    assert rc == L.ERR_STREAM and "pclr" in L.last_error()
