"""Host Tier-2 writer (packets, tag trees, layer formation, markers) checked byte for byte against the
compiled reference, with the Tier-1 results supplied by the ORACLE (no GPU needed): the whole-frame
codestream must equal ref_encode's output for identical parameters."""
import numpy as np
import pytest
import hostpipe
import refgrok
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame

pytestmark = pytest.mark.skipif(not refgrok.available(), reason="oracle/_ref not built (needs /root/reference)")

CASES = [
    # h, w, ncomp, prec, kind, mct, numres, cblk, tile, prog
    (96, 128, 3, 8, "smooth", 1, 6, 64, 0, 0),
    (96, 128, 3, 8, "noise", 0, 6, 64, 0, 0),
    (75, 133, 1, 8, "ramp", 0, 4, 32, 0, 1),
    (100, 140, 4, 16, "smooth", 1, 5, 64, 0, 0),
    (130, 200, 3, 16, "noise", 1, 3, 32, 64, 2),
    (90, 150, 3, 12, "smooth", 1, 6, 64, 0, 4),
    (64, 64, 3, 8, "ramp", 1, 1, 64, 0, 3),
]


@pytest.mark.parametrize("h,w,ncomp,prec,kind,mct,numres,cblk,tile,prog", CASES)
def test_lossless_codestream_identical(h, w, ncomp, prec, kind, mct, numres, cblk, tile, prog):
    img = make_frame(h, w, ncomp, prec, kind, seed=1)
    planes = [img[..., c].astype(np.int32) for c in range(ncomp)]
    p = L.default_params(width=w, height=h, ncomp=ncomp, prec=prec, mct=mct, numresolution=numres,
                         cblk_w=cblk, cblk_h=cblk, tile_w=tile, tile_h=tile, prog_order=prog)
    ours = hostpipe.encode(planes, prec, p)
    ref = refgrok.encode(planes, prec, mct=mct, numresolution=numres, cblk=(cblk, cblk), tile=(tile, tile), prog=prog)
    assert len(ours) == len(ref)
    assert ours == ref
    dec = refgrok.decode(ours)
    for a, b in zip(dec, planes):
        assert np.array_equal(a, b)
