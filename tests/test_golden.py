"""Known answers: tests/golden/*.npz were written by the reference itself (ext/grok compiled from /root/reference) and by
OpenJPEG 2.5.x (tests/golden/make_golden.py, run in the build container).  CPU: the oracle and the product's host Tier-2
reproduce them.  GPU: the CUDA path reproduces them through the C ABI -- also on a box where /root/reference does not exist."""
import ctypes as C
import os
import numpy as np
import pytest
import orc
import hostpipe
from j2k_b200 import _lib as L

HERE = os.path.dirname(os.path.abspath(__file__))
ST = np.load(os.path.join(HERE, "golden", "stages.npz"))
CS = np.load(os.path.join(HERE, "golden", "codestreams.npz"))
LOSSLESS = {"c1_rgb8": dict(prec=8, mct=1), "c2_rgba16": dict(prec=16, mct=1), "c4_tiles": dict(prec=16, mct=1, tile_w=64, tile_h=64, numresolution=4),
            "plugin_as_written": dict(prec=8, mct=0, numlayers=12, tile_w=64, tile_h=64, numresolution=4)}


def test_oracle_stages_reproduce_the_reference_vectors():
    c = list(ST["mct_in"])
    assert np.array_equal(np.stack(orc.mct(*c)), ST["rct_fwd"])
    assert np.array_equal(np.stack(orc.mct(*c, inverse=True)), ST["rct_inv"])
    assert np.array_equal(np.stack(orc.mct(*[x * 2048 for x in c], real=True)), ST["ict_fwd"])
    assert np.array_equal(np.stack(orc.mct(*list(ST["ict_inv_in"]), inverse=True, real=True)).view(np.int32), ST["ict_inv"].view(np.int32))
    assert np.array_equal(orc.dwt(ST["dwt_in"], 4, x0=3, y0=5), ST["dwt53_fwd"])
    assert np.array_equal(orc.dwt(ST["dwt_in"], 4, inverse=True, x0=3, y0=5), ST["dwt53_inv"])
    assert np.array_equal(orc.dwt(ST["dwt_in"] * 8, 4, real=True), ST["dwt97_fwd"])
    assert np.array_equal(orc.dwt(ST["dwt97_inv_in"], 4, inverse=True, real=True).view(np.int32), ST["dwt97_inv"].view(np.int32))


@pytest.mark.parametrize("i", range(5))
def test_oracle_tier1_reproduces_the_reference_blocks(i):
    b = ST["blk%d_in" % i]
    orient, numbps, npasses, cut = [int(v) for v in ST["blk%d_meta" % i]]
    h, w = b.shape
    q, mx = orc.quantize(b)
    r = orc.t1_encode(q, orient, mx)
    assert (r["numbps"], r["npasses"]) == (numbps, npasses) and r["rates"] == list(ST["blk%d_rates" % i])
    assert r["data"] == ST["blk%d_bytes" % i].tobytes()
    dist = orc.distortions(r["nmsedec"], numbps, 1.0, orc.lib().orc_dwt_norm(0, orient, 0), 1.0)
    assert np.allclose(dist, ST["blk%d_dist" % i], rtol=1e-12, atol=0)
    assert np.array_equal(orc.t1_decode(r["data"], w, h, orient, numbps, npasses), ST["blk%d_dec" % i])
    rc = r["rates"][cut - 1]
    assert np.array_equal(orc.t1_decode(r["data"][:rc], w, h, orient, numbps, cut), ST["blk%d_dec_cut" % i])


@pytest.mark.parametrize("name", sorted(LOSSLESS))
def test_host_tier2_reproduces_the_reference_codestreams(name):
    img = CS[name + "_img"]
    kw = dict(LOSSLESS[name]); prec = kw.pop("prec")
    h, w, nc = img.shape
    p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, **kw)
    assert hostpipe.encode([img[..., c].astype(np.int32) for c in range(nc)], prec, p) == CS[name + "_stream"].tobytes()


def test_host_tier2_reproduces_openjpeg_2_5_streams():
    img = CS["opj254_rgb8_img"]
    p = L.default_params(width=img.shape[1], height=img.shape[0], ncomp=3, prec=8, mct=1, compat=L.COMPAT_OPENJPEG)
    assert hostpipe.encode([img[..., c].astype(np.int32) for c in range(3)], 8, p) == CS["opj254_rgb8_stream"].tobytes()


# ---- the CUDA path against the same fixtures ----
@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(LOSSLESS))
def test_cuda_encode_and_decode_reproduce_the_reference_codestreams(name):
    import j2k
    img = CS[name + "_img"]
    kw = dict(LOSSLESS[name]); prec = kw.pop("prec")
    ref = CS[name + "_stream"].tobytes()
    assert j2k.encode(np.ascontiguousarray(img), prec, **kw) == ref
    assert np.array_equal(j2k.decode(ref), img)


@pytest.mark.gpu
def test_cuda_cinema_and_rate_capped_fixtures():
    import j2k
    ref = CS["c3_cinema2k_stream"].tobytes()
    dec = j2k.decode(ref)                                            # 9/7 + ICT stream written by the reference: +-1 LSB of its decoder
    assert np.abs(dec[0].astype(np.int32) - CS["c3_cinema2k_row0"].astype(np.int32)).max() <= 1
    s = int(dec.astype(np.int64).sum())
    assert abs(s - int(CS["c3_cinema2k_pixsum"][0])) <= dec.size // 50      # rounding differences average out far below one LSB per sample
    img = CS["layers_rate_img"]
    ours = j2k.encode(np.ascontiguousarray(img), 8, irreversible=1, numlayers=3, rates=(20, 8, 3))
    assert ours == CS["layers_rate_stream"].tobytes()                # three rate-capped layers: PCRD byte for byte
    got = j2k.decode(ours)
    assert np.abs(got.astype(np.int32) - CS["layers_rate_decoded"].astype(np.int32)).max() <= 1


@pytest.mark.gpu
def test_cuda_stage_kernels_reproduce_the_reference_vectors():
    import j2k
    c = list(ST["mct_in"])
    assert np.array_equal(np.stack(j2k.stage_mct(*c)), ST["rct_fwd"])
    assert np.array_equal(np.stack(j2k.stage_mct(*c, inverse=1)), ST["rct_inv"])
    assert np.array_equal(np.stack(j2k.stage_mct(*[x * 2048 for x in c], irreversible=1)), ST["ict_fwd"])
    assert np.array_equal(np.stack(j2k.stage_mct(*list(ST["ict_inv_in"]), irreversible=1, inverse=1)).view(np.int32), ST["ict_inv"].view(np.int32))
    assert np.array_equal(j2k.stage_dwt(ST["dwt_in"], 4, x0=3, y0=5), ST["dwt53_fwd"])
    assert np.array_equal(j2k.stage_dwt(ST["dwt_in"], 4, inverse=1, x0=3, y0=5), ST["dwt53_inv"])
    assert np.array_equal(j2k.stage_dwt(ST["dwt_in"] * 8, 4, irreversible=1), ST["dwt97_fwd"])
    assert np.array_equal(j2k.stage_dwt(ST["dwt97_inv_in"], 4, irreversible=1, inverse=1).view(np.int32), ST["dwt97_inv"].view(np.int32))
    for i in range(5):
        b = ST["blk%d_in" % i]
        orient, numbps, npasses, cut = [int(v) for v in ST["blk%d_meta" % i]]
        h, w = b.shape
        r = j2k.stage_t1_encode(b, [(0, 0, w, h, orient, 0)])[0]
        assert (r["numbps"], r["npasses"], r["rates"]) == (numbps, npasses, list(ST["blk%d_rates" % i]))
        assert r["data"] == ST["blk%d_bytes" % i].tobytes()
        rc = int(ST["blk%d_rates" % i][cut - 1])
        dec = j2k.stage_t1_decode([(w, h, orient, numbps, npasses, r["data"]), (w, h, orient, numbps, cut, r["data"][:rc])])
        assert np.array_equal(dec[0], ST["blk%d_dec" % i]) and np.array_equal(dec[1], ST["blk%d_dec_cut" % i])
