"""GPU decode parity through the C ABI: Tier-1 decode vs the oracle per block (reference decoder
representation), whole-frame decode bit-exact vs the source pixels and the compiled reference's decoder,
reduced-resolution decode, tiles, multi-layer streams, 9/7 within +-1 LSB of the reference decoder."""
import numpy as np
import pytest
import orc
import j2k
import refgrok
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("scale", [2, 40, 3000, 60000])
def test_t1_decode_blocks(scale):
    rng = np.random.default_rng(scale)
    shapes = [(64, 64), (32, 32), (17, 9), (64, 3), (5, 64), (1, 1), (33, 37), (64, 30)]
    blocks, want = [], []
    for i, (w, h) in enumerate(shapes):
        c = rng.laplace(0, scale, size=(h, w)).astype(np.int32)
        if scale == 2:
            c[rng.random((h, w)) < 0.8] = 0
        q, mx = orc.quantize(c)
        r = orc.t1_encode(q, i % 4, mx)
        for npasses in sorted(set([r["npasses"], max(r["npasses"] - 1, 0), max(r["npasses"] - 2, 0), r["npasses"] // 2])):
            if npasses == 0:
                continue
            data = r["data"][:r["rates"][npasses - 1]]
            blocks.append((w, h, i % 4, r["numbps"], npasses, data))
            want.append(orc.t1_decode(data, w, h, i % 4, r["numbps"], npasses))
    got = j2k.stage_t1_decode(blocks)
    for b, g, wnt in zip(blocks, got, want):
        d = np.argwhere(g != wnt)
        assert len(d) == 0, "block %dx%d orient %d numbps %d passes %d: %d diffs, first %s got %d want %d" % (
            b[0], b[1], b[2], b[3], b[4], len(d), d[0], g[tuple(d[0])], wnt[tuple(d[0])])


CASES = [
    (96, 128, 3, 8, "smooth", 1, 6, 64, 0, 0),
    (270, 480, 3, 8, "noise", 0, 6, 64, 0, 0),
    (75, 133, 1, 8, "ramp", 0, 4, 32, 0, 1),
    (200, 340, 4, 16, "smooth", 1, 5, 64, 0, 0),
    (130, 200, 3, 16, "noise", 1, 3, 32, 64, 2),
    (190, 350, 3, 12, "smooth", 1, 6, 64, 0, 4),
    (432, 820, 3, 16, "smooth", 1, 6, 64, 256, 0),
    (64, 64, 3, 8, "ramp", 1, 1, 64, 0, 3),
]


@pytest.mark.parametrize("h,w,ncomp,prec,kind,mct,numres,cblk,tile,prog", CASES)
def test_lossless_roundtrip_and_reference_streams(h, w, ncomp, prec, kind, mct, numres, cblk, tile, prog):
    img = make_frame(h, w, ncomp, prec, kind, seed=5)
    ours = j2k.encode(img, prec, mct=mct, numresolution=numres, cblk_w=cblk, cblk_h=cblk, tile_w=tile, tile_h=tile, prog_order=prog)
    assert np.array_equal(j2k.decode(ours), img)
    if refgrok.available():
        planes = [img[..., c].astype(np.int32) for c in range(ncomp)]
        ref = refgrok.encode(planes, prec, mct=mct, numresolution=numres, cblk=(cblk, cblk), tile=(tile, tile), prog=prog)
        assert np.array_equal(j2k.decode(ref), img)


@pytest.mark.parametrize("reduce", [1, 2, 3])
@pytest.mark.parametrize("tile", [0, 128])
def test_reduced_resolution_matches_reference(reduce, tile):
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    img = make_frame(300, 421, 3, 8, "smooth", seed=6)
    cs = j2k.encode(img, 8, mct=1, numresolution=5, tile_w=tile, tile_h=tile)
    got = j2k.decode(cs, reduce=reduce)
    ref = refgrok.decode(cs, reduce=reduce)
    for c in range(3):
        assert np.array_equal(got[..., c].astype(np.int32), ref[c])


def test_multilayer_rate_capped_reference_stream():
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    img = make_frame(256, 384, 3, 8, "smooth", seed=7)
    planes = [img[..., c].astype(np.int32) for c in range(3)]
    cs = refgrok.encode(planes, 8, mct=1, numlayers=3, rates=[60.0, 20.0, 8.0])     # lossy last layer: truncated passes
    got = j2k.decode(cs)
    ref = refgrok.decode(cs)
    for c in range(3):
        assert np.array_equal(got[..., c].astype(np.int32), ref[c])


@pytest.mark.parametrize("prec,ncomp", [(8, 3), (12, 3), (16, 1)])
def test_irreversible_97_within_one_lsb_of_reference_decoder(prec, ncomp):
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    img = make_frame(240, 352, ncomp, prec, "smooth", seed=8)
    planes = [img[..., c].astype(np.int32) for c in range(ncomp)]
    cs = refgrok.encode(planes, prec, irreversible=1, mct=1 if ncomp == 3 else 0, rates=[10.0])
    got = j2k.decode(cs)
    ref = refgrok.decode(cs)
    for c in range(ncomp):
        diff = np.abs(got[..., c].astype(np.int64) - ref[c])
        assert diff.max() <= 1, "component %d: max diff %d" % (c, diff.max())      # north_star tolerance: +-1 LSB


def test_4k_rgba16_roundtrip():
    """BASELINE config 2 at full size: encode -> decode is the identity"""
    img = make_frame(2160, 4096, 4, 16, "smooth", seed=9)
    cs = j2k.encode(img, 16, mct=1)
    assert np.array_equal(j2k.decode(cs), img)


# ---- decode windows (row f3, second half): opj_set_decode_area ----
@pytest.mark.gpu
@pytest.mark.parametrize("irreversible,tile,reduce", [(0, 0, 0), (1, 0, 0), (0, 128, 0), (0, 0, 1), (1, 128, 2)])
@pytest.mark.parametrize("win", [(0, 0, 64, 48), (301, 97, 420, 203), (250, 0, 512, 30), (100, 300, 101, 301)])
def test_decode_window_equals_the_crop_of_the_full_decode(irreversible, tile, reduce, win):
    import ctypes as C
    img = make_frame(384, 512, 3, 8, "noise", seed=17)
    cs = j2k.encode(img, 8, mct=1, irreversible=irreversible, tile_w=tile, tile_h=tile, numresolution=5)
    full = j2k.decode(cs, reduce=reduce)
    x0, y0, x1, y1 = win
    f = 1 << reduce
    rx0, ry0, rx1, ry1 = -(-x0 // f), -(-y0 // f), -(-x1 // f), -(-y1 // f)
    if rx1 <= rx0 or ry1 <= ry0:
        return
    out = np.zeros((ry1 - ry0, rx1 - rx0, 3), np.uint8)
    lib = L.lib()
    d = C.c_void_p(lib.j2kcuda_decoder_create(-1))
    assert lib.j2kcuda_decoder_set_window(d, x0, y0, x1, y1) == 0
    assert lib.j2kcuda_decoder_prepare(d, cs, C.c_size_t(len(cs)), reduce) == 0, L.last_error()
    assert lib.j2kcuda_decoder_run(d) == 0, L.last_error()
    assert lib.j2kcuda_decoder_download(d, j2k.planes_of(out), 3) == 0, L.last_error()
    cnt = (C.c_uint64 * 6)()
    lib.j2kcuda_decoder_counters(d, cnt, 6)
    lib.j2kcuda_decoder_destroy(d)
    assert np.array_equal(out, full[ry0:ry1, rx0:rx1])
    if (x1 - x0) * (y1 - y0) < 20000:
        assert cnt[2] > 0                                            # a small window skips Tier-1 work
    out2 = np.zeros_like(out)
    assert lib.j2kcuda_decode_region(cs, C.c_size_t(len(cs)), reduce, x0, y0, x1, y1, j2k.planes_of(out2), 3) == 0, L.last_error()
    assert np.array_equal(out2, out)
