"""Row a21: ENCODING with the code-block styles RESET (0x02), TERMALL (0x04) and SEGSYM (0x20) -- styles the arithmetic-coder
stage can do alone.  Refused: LAZY (needs raw sign bits), VSC (changes the context modelling), PTERM (the reference's own PTERM
streams of small images decode neither in its decoder nor in OpenJPEG, and there is nothing to pin a corrected one against).

CPU: oracle stages + the product's Tier-2 writer (tests/hostpipe.py) against the compiled reference.  RESET streams are the
reference's byte for byte.  TERMALL streams are the reference's byte for byte wherever the reference's own stream can be decoded;
where it cannot (a terminated pass of exactly one byte gets length 0, t1.cpp:1649-1651 -- the reference's decoder and OpenJPEG
2.5.4 both fail on the reference's stream) ours differs and is read correctly by both.
GPU: the CUDA path (k_t1_context + k_t1_mq_gen) writes the same bytes as the CPU pipeline, and decodes them again."""
import io
import numpy as np
import pytest
import hostpipe
import refgrok
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame

# (h, w, ncomp, prec, parameters) -- sizes the reference's TERMALL encoder survives (it corrupts its heap on small code-blocks and
# on 12/16-bit input: another reason not to copy its buffer arithmetic)
CASES = [
    (96, 128, 3, 8, dict(mct=1)),
    (70, 53, 1, 12, dict(numresolution=3)),
    (128, 128, 3, 8, dict(mct=1, irreversible=1, numlayers=3, rates=(40, 20, 8))),
    (100, 150, 3, 8, dict(mct=1, tile_w=64, tile_h=64, numresolution=4, numlayers=2, rates=(20, 0))),
]


def planes_of(img):
    return [img[:, :, c].astype(np.int32) for c in range(img.shape[2])]


def ref_encode(planes, prec, sty, kw):
    refgrok.set_extra(mode=sty)
    try:
        return refgrok.encode(planes, prec, irreversible=kw.get("irreversible", 0), numresolution=kw.get("numresolution", 6),
                              mct=kw.get("mct", 0), tile=(kw.get("tile_w", 0), kw.get("tile_h", 0)), numlayers=kw.get("numlayers", 1),
                              rates=kw.get("rates", ()))
    finally:
        refgrok.set_extra()


def lossless(kw):
    """every pass ends up in the stream: 5/3 and a last layer without a rate"""
    r = kw.get("rates", ())
    return not kw.get("irreversible") and (not r or r[-1] == 0)


STYLES = [0x02, 0x04, 0x06, 0x20, 0x26]      # RESET 2, TERMALL 4, SEGSYM 0x20


@pytest.mark.parametrize("case", range(len(CASES)))
@pytest.mark.parametrize("sty", STYLES)
def test_host_pipeline_against_the_reference(sty, case):
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    h, w, nc, prec, kw = CASES[case]
    img = make_frame(h, w, nc, prec, "smooth", seed=sty + h)
    planes = planes_of(img)
    ours = hostpipe.encode(planes, prec, L.default_params(width=w, height=h, ncomp=nc, prec=prec, cblksty=sty, **kw))
    ref = ref_encode(planes, prec, sty, kw)
    back = refgrok.decode(ours)
    if lossless(kw):
        assert all(np.array_equal(back[c], planes[c]) for c in range(nc)), "the reference's decoder cannot read our stream"
    if ours != ref:
        assert sty & 4, "without TERMALL the stream must be the reference's"
        assert lossless(kw)
        theirs = refgrok.decode(ref)
        assert not all(np.array_equal(theirs[c], planes[c]) for c in range(nc)), "streams differ although the reference's own decodes"
    if lossless(kw) and prec == 8:
        from PIL import Image                                  # libopenjp2 2.5.4
        im = np.array(Image.open(io.BytesIO(ours)))
        assert np.array_equal(im.reshape(h, w, nc), img)


@pytest.mark.parametrize("sty", [2, 4, 6, 0x20, 0x26])
def test_small_images_whose_deep_levels_hit_the_reference_defect(sty):
    """64 x 64 with six resolutions: 2 x 2 code-blocks at the top of the pyramid, terminated passes of one byte"""
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    from PIL import Image
    img = make_frame(64, 64, 3, 8, "smooth", seed=sty)
    ours = hostpipe.encode(planes_of(img), 8, L.default_params(width=64, height=64, ncomp=3, prec=8, mct=1, cblksty=sty))
    back = refgrok.decode(ours)
    assert all(np.array_equal(back[c], img[:, :, c]) for c in range(3))
    assert np.array_equal(np.array(Image.open(io.BytesIO(ours))), img)


def test_other_style_bits_are_refused():
    for sty in (1, 8, 0x10, 7, 0x40):
        p = L.default_params(width=64, height=64, ncomp=1, prec=8, cblksty=sty)
        assert L.lib().j2kcuda_t2_block_count(L.C.byref(p) if hasattr(L, "C") else __import__("ctypes").byref(p)) < 0
        assert "RESET" in L.last_error()


@pytest.mark.gpu
@pytest.mark.parametrize("sty", STYLES)
def test_cuda_path_writes_the_cpu_pipelines_bytes_and_reads_them_back(sty):
    import j2k
    for h, w, nc, prec, kw in CASES + [(64, 64, 3, 8, dict(mct=1)), (270, 480, 3, 16, dict(mct=1)), (33, 47, 1, 10, dict(cblk_w=16, cblk_h=16, numresolution=4))]:
        img = make_frame(h, w, nc, prec, "smooth", seed=sty + h)
        cs = j2k.encode(img, prec, cblksty=sty, **kw)
        want = hostpipe.encode(planes_of(img), prec, L.default_params(width=w, height=h, ncomp=nc, prec=prec, cblksty=sty, **kw))
        assert cs == want, (sty, h, w, nc, prec, kw)
        back = j2k.decode(cs)
        if lossless(kw):
            assert np.array_equal(back, img)
