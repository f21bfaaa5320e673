"""Whole-frame encode through the reference-facing C-ABI call (j2kcuda_encode, host buffers in the After
Effects interleaved layout): lossless codestreams byte-identical to the compiled reference for identical
parameters, decoded by the reference's own decompressor bit-exact."""
import numpy as np
import pytest
import j2k
import refgrok
from j2k_b200.frames import make_frame

pytestmark = pytest.mark.gpu

CASES = [
    (96, 128, 3, 8, "smooth", 1, 6, 64, 0, 0),
    (270, 480, 3, 8, "noise", 0, 6, 64, 0, 0),
    (75, 133, 1, 8, "ramp", 0, 4, 32, 0, 1),
    (200, 340, 4, 16, "smooth", 1, 5, 64, 0, 0),
    (130, 200, 3, 16, "noise", 1, 3, 32, 64, 2),
    (190, 350, 3, 12, "smooth", 1, 6, 64, 0, 4),
    (540, 1024, 3, 8, "smooth", 1, 6, 64, 0, 0),
    (432, 820, 3, 16, "smooth", 1, 6, 64, 256, 0),
]


@pytest.mark.parametrize("h,w,ncomp,prec,kind,mct,numres,cblk,tile,prog", CASES)
def test_lossless_identical_to_reference(h, w, ncomp, prec, kind, mct, numres, cblk, tile, prog):
    img = make_frame(h, w, ncomp, prec, kind, seed=2)
    ours = j2k.encode(img, prec, mct=mct, numresolution=numres, cblk_w=cblk, cblk_h=cblk, tile_w=tile, tile_h=tile,
                      prog_order=prog)
    planes = [img[..., c].astype(np.int32) for c in range(ncomp)]
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    ref = refgrok.encode(planes, prec, mct=mct, numresolution=numres, cblk=(cblk, cblk), tile=(tile, tile), prog=prog)
    assert len(ours) == len(ref), (len(ours), len(ref))
    assert ours == ref
    for a, b in zip(refgrok.decode(ours), planes):
        assert np.array_equal(a, b)


def test_config1_2k_rgb8():
    """BASELINE config 1: 2048x1080 8-bit RGB lossless 5/3, 5 levels, 64x64, single tile"""
    img = make_frame(1080, 2048, 3, 8, "smooth", seed=0)
    ours = j2k.encode(img, 8, mct=1, numresolution=6)
    if not refgrok.available():
        pytest.skip("oracle/_ref not built")
    planes = [img[..., c].astype(np.int32) for c in range(3)]
    assert ours == refgrok.encode(planes, 8, mct=1, numresolution=6, threads=8)
