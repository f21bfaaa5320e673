"""Host Tier-2 reader (no GPU): parsing a codestream written by the reference must give back, per code-block,
the zero-bit-plane count, pass count and bytes the encoder side put in; main-header probe fields."""
import numpy as np
import pytest
import hostpipe
import orc
import refgrok
import j2k
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame

pytestmark = pytest.mark.skipif(not refgrok.available(), reason="oracle/_ref not built (needs /root/reference)")

CASES = [(96, 128, 3, 8, 1, 6, 64, 0, 0, 1), (100, 140, 4, 16, 1, 5, 64, 0, 0, 1), (130, 200, 3, 16, 1, 3, 32, 64, 2, 1),
         (90, 150, 3, 12, 1, 6, 64, 0, 4, 1), (120, 160, 3, 8, 1, 5, 32, 0, 1, 3), (64, 64, 1, 8, 0, 1, 64, 0, 3, 1)]


@pytest.mark.parametrize("h,w,ncomp,prec,mct,numres,cblk,tile,prog,layers", CASES)
def test_parse_recovers_block_streams(h, w, ncomp, prec, mct, numres, cblk, tile, prog, layers):
    img = make_frame(h, w, ncomp, prec, "smooth", seed=3)
    planes = [img[..., c].astype(np.int32) for c in range(ncomp)]
    ref = refgrok.encode(planes, prec, mct=mct, numresolution=numres, cblk=(cblk, cblk), tile=(tile, tile), prog=prog,
                         numlayers=layers, rates=[40.0, 10.0, 0.0][:layers] if layers > 1 else ())
    info = j2k.probe(ref)
    assert (info.width, info.height, info.ncomp, info.prec[0]) == (w, h, ncomp, prec)
    assert info.reversible == 1 and info.numresolution == numres and info.numlayers == layers and info.mct == mct
    assert (info.cblk_w, info.cblk_h, info.prog_order) == (cblk, cblk, prog)
    got = j2k.t2_parse(ref)
    p = L.default_params(width=w, height=h, ncomp=ncomp, prec=prec, mct=mct, numresolution=numres, cblk_w=cblk,
                         cblk_h=cblk, tile_w=tile, tile_h=tile, prog_order=prog)
    geo = hostpipe.block_list(p)
    assert len(got) == len(geo)
    tw, th = (tile or w), (tile or h)
    ntx = -(-w // tw)
    coeffs = {}
    checked = 0
    for g, (numbps, npasses, data) in zip(geo, got):
        t, comp, px, py, bw, bh, orient, level, band_numbps = [int(v) for v in g[:9]]
        if bw == 0 or bh == 0:
            assert npasses == 0
            continue
        if t not in coeffs:
            tx, ty = t % ntx, t // ntx
            coeffs[t] = hostpipe.oracle_tile_stages([pl[ty * th:(ty + 1) * th, tx * tw:(tx + 1) * tw] for pl in planes], prec, 0, p)
        q, mx = orc.quantize(coeffs[t][comp][py:py + bh, px:px + bw])
        want = orc.t1_encode(q, orient, mx)
        if layers == 1:
            # grok drops trailing passes that add neither rate nor distortion (tcd.cpp:373-377)
            assert npasses <= want["npasses"]
            assert data == want["data"][:len(data)]
        if npasses:
            assert numbps == want["numbps"]
            out = orc.t1_decode(data, bw, bh, orient, numbps, npasses)
            # all layers parsed and the last layer is lossless, so the block decodes exactly
            if True:
                assert np.array_equal(np.fix(out / 2).astype(np.int32), coeffs[t][comp][py:py + bh, px:px + bw])
            checked += 1
    assert checked > 0


def test_rejects_garbage():
    import ctypes as C
    info = L.Info()
    assert L.lib().j2kcuda_probe(b"not a codestream", C.c_size_t(16), C.byref(info)) == L.ERR_STREAM
