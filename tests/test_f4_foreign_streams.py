"""Row f4: what third-party codestreams use and the encoder of this library never writes -- code-block styles (LAZY / bypass, RESET,
TERMALL, VSC, PTERM, SEGSYM), several codeword segments per block, the ROI up-shift, SOP / EPH markers, sub-sampled components.
The streams come from the reference encoder (ext/grok compiled into oracle/_ref, `params.mode` etc. through ref_shim);
pixels are compared with the reference decoder's, bit-exact for 5/3."""
import ctypes as C
import numpy as np
import pytest
import refgrok
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame

pytestmark = pytest.mark.skipif(not refgrok.available(), reason="oracle/_ref not built (needs /root/reference)")
LAZY, RESET, TERMALL, VSC, PTERM, SEGSYM = 1, 2, 4, 8, 16, 32
STYLES = [LAZY, RESET, TERMALL, VSC, PTERM, SEGSYM, LAZY | TERMALL, RESET | VSC | SEGSYM, 63]


def ref_stream(img, prec, **extra):
    kw = {k: extra.pop(k) for k in list(extra) if k in ("mode", "roi_comp", "roi_shift", "csty", "sub")}
    refgrok.set_extra(**kw)
    try:
        nc = img.shape[2]
        sub = kw.get("sub", (1, 1))
        planes = [img[..., 0].astype(np.int32)] + [np.ascontiguousarray(img[::sub[1], ::sub[0], c]).astype(np.int32) for c in range(1, nc)]
        return refgrok.encode(planes, prec, **extra)
    finally:
        refgrok.set_extra()


@pytest.mark.parametrize("sty", STYLES)
def test_parser_reads_the_segments_of_every_code_block_style(sty):
    import j2k
    img = make_frame(70, 90, 3, 8, "noise", seed=sty)
    cs = ref_stream(img, 8, mode=sty, mct=1, numlayers=2, rates=(8, 0), irreversible=0)
    blocks = j2k.t2_parse(cs)                                        # host only
    assert sum(len(b[2]) for b in blocks) > 1000 and all(b[1] <= 3 * b[0] - 2 for b in blocks if b[0])


@pytest.mark.gpu
@pytest.mark.parametrize("sty", STYLES)
@pytest.mark.parametrize("irreversible", [0, 1])
def test_code_block_styles_decode_like_the_reference(sty, irreversible):
    import j2k
    img = make_frame(150, 200, 3, 12, "smooth" if sty & 1 else "noise", seed=sty + irreversible).astype(np.uint16)
    kw = dict(mode=sty, mct=1, irreversible=irreversible, cblk=(32, 32) if sty & 4 else (64, 64))
    if irreversible:
        kw.update(numlayers=2, rates=(20, 4))
    cs = ref_stream(img, 12, **kw)
    want = np.stack(refgrok.decode(cs), axis=-1)
    got = j2k.decode(cs).astype(np.int32)
    if irreversible:
        assert np.abs(got - want).max() <= 1
    else:
        assert np.array_equal(got, want)
        if not (sty & LAZY):                      # (the reference's own bypass ENCODER does not round-trip: its decoder returns other pixels too)
            assert np.array_equal(got, img)


@pytest.mark.gpu
@pytest.mark.parametrize("csty", [2, 4, 6])
def test_sop_and_eph_markers(csty):
    import j2k
    img = make_frame(128, 160, 3, 8, "smooth", seed=csty)          # (tiles at least 2^levels wide: see DESIGN.md on the reference's phantom blocks)
    cs = ref_stream(img, 8, csty=csty, mct=1, tile=(64, 64), numresolution=4)
    assert np.array_equal(np.stack(refgrok.decode(cs), axis=-1), img)
    assert np.array_equal(j2k.decode(cs), img)


@pytest.mark.gpu
@pytest.mark.parametrize("shift", [3, 9])
def test_region_of_interest_shift(shift):
    import j2k
    img = make_frame(96, 128, 3, 8, "smooth", seed=shift)
    cs = ref_stream(img, 8, roi_comp=1, roi_shift=shift, mct=0)
    want = np.stack(refgrok.decode(cs), axis=-1)
    assert np.array_equal(j2k.decode(cs).astype(np.int32), want)


@pytest.mark.gpu
@pytest.mark.parametrize("sub", [(2, 2), (2, 1), (1, 2), (3, 2)])
@pytest.mark.parametrize("reduce", [0, 1])
def test_sub_sampled_components_are_replicated_like_copybuffer(sub, reduce):
    """YCbCr 4:2:0 / 4:2:2 style files: components at their own size in the stream; the output replicates them to the reference
    grid the way Codec::CopyBuffer does above the reference codec (src/common/j2k_codec.cpp:262-279)"""
    import j2k
    img = make_frame(97, 131, 3, 8, "smooth", seed=sub[0] * 4 + sub[1])
    cs = ref_stream(img, 8, sub=sub, mct=0)
    comps = refgrok.decode_components(cs, reduce=reduce)
    h, w = comps[0].shape
    got = j2k.decode(cs, reduce=reduce)
    assert got.shape[:2] == (h, w) and np.array_equal(got[..., 0], comps[0])
    for c in (1, 2):
        rep = comps[c][np.minimum(np.arange(h) // sub[1], comps[c].shape[0] - 1)][:, np.minimum(np.arange(w) // sub[0], comps[c].shape[1] - 1)]
        assert np.array_equal(got[..., c], rep)
