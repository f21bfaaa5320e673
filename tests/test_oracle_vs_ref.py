"""Pins the plain-C oracle (oracle/j2k_oracle.c) against the reference's own sources compiled here
(oracle/_ref/libgrokref.so): stage arrays exact, per-block Tier-1 results exact."""
import numpy as np
import pytest
import orc
import refgrok

pytestmark = pytest.mark.skipif(not refgrok.available(), reason="oracle/_ref not built (needs /root/reference)")


def rnd(shape, lo, hi, seed):
    return np.random.default_rng(seed).integers(lo, hi, size=shape, dtype=np.int64).astype(np.int32)


@pytest.mark.parametrize("real", [False, True])
def test_mct_forward(real):
    c = [rnd((4099,), -70000 if real else -40000, 70000 if real else 40000, s) for s in range(3)]
    if real:
        c = [x * 2048 // 64 for x in c]
    a = orc.mct(*c, real=real)
    b = refgrok.mct(*c, real=real)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_rct_inverse():
    c = [rnd((1001,), -40000, 40000, s) for s in range(3)]
    for x, y in zip(orc.mct(*c, inverse=True), refgrok.mct(*c, inverse=True)):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("shape,x0,y0,numres", [((64, 64), 0, 0, 6), ((37, 53), 0, 0, 4), ((33, 17), 3, 5, 3),
                                                ((1, 40), 0, 0, 3), ((40, 1), 1, 1, 3), ((130, 257), 0, 0, 6)])
@pytest.mark.parametrize("real", [False, True])
def test_dwt_forward(shape, x0, y0, numres, real):
    a = rnd(shape, -40000, 40000, 7) * (8 if real else 1)
    assert np.array_equal(orc.dwt(a, numres, real=real, x0=x0, y0=y0), refgrok.dwt_encode(a, numres, real=int(real), x0=x0, y0=y0))


@pytest.mark.parametrize("shape,x0,y0,numres", [((64, 64), 0, 0, 6), ((37, 53), 0, 0, 4), ((33, 17), 3, 5, 3), ((130, 257), 0, 0, 6)])
def test_dwt53_inverse(shape, x0, y0, numres):
    a = rnd(shape, -40000, 40000, 9)
    assert np.array_equal(orc.dwt(a, numres, inverse=True, x0=x0, y0=y0), refgrok.dwt_decode(a, numres, x0=x0, y0=y0))


@pytest.mark.parametrize("shape,numres", [((64, 64), 6), ((37, 53), 4), ((130, 257), 6)])
def test_dwt97_inverse_float(shape, numres):
    a = np.random.default_rng(3).normal(0, 500, size=shape).astype(np.float32)
    got = orc.dwt(a, numres, inverse=True, real=True)
    ref = refgrok.dwt_decode(a.view(np.int32), numres, real=1).view(np.float32)
    assert np.allclose(got, ref, rtol=0, atol=2e-3 * np.abs(ref).max())


@pytest.mark.parametrize("w,h", [(64, 64), (32, 32), (17, 9), (64, 3), (5, 64), (1, 1)])
@pytest.mark.parametrize("orient", [0, 1, 2, 3])
@pytest.mark.parametrize("scale", [3, 300, 30000])
def test_t1_block(w, h, orient, scale):
    rng = np.random.default_rng(w * 131 + h * 7 + orient + scale)
    c = (rng.laplace(0, scale, size=(h, w))).astype(np.int32)
    if scale == 3:
        c[rng.random((h, w)) < 0.7] = 0
    ref = refgrok.t1_encode_cblk(c, orient)
    q, mx = orc.quantize(c)
    got = orc.t1_encode(q, orient, mx)
    assert got["numbps"] == ref["numbps"] and got["npasses"] == ref["npasses"]
    assert got["rates"] == ref["rates"]
    assert got["data"] == ref["data"]
    dist = orc.distortions(got["nmsedec"], got["numbps"], 1.0, orc.lib().orc_dwt_norm(0, orient, 0), 1.0)
    assert np.allclose(dist, ref["dist"], rtol=1e-12, atol=0)


def test_t1_roundtrip_decode():
    rng = np.random.default_rng(5)
    c = rng.laplace(0, 200, size=(48, 64)).astype(np.int32)
    q, mx = orc.quantize(c)
    r = orc.t1_encode(q, 1, mx)
    out = orc.t1_decode(r["data"], 64, 48, 1, r["numbps"], r["npasses"])
    assert np.array_equal(np.fix(out / 2).astype(np.int32), c)     # T1Decoder.cpp:77-99: C truncating division


# ---- stage-level pins the round-1 review asked for: inverse ICT, the Tier-1 DECODER, and a tight 9/7 inverse ----
def test_ict_inverse_float_is_exact():
    """orc_ict_inv against opj_mct_decode_real (mct.cpp:406-471) directly, bit for bit"""
    rng = np.random.default_rng(11)
    c = [rng.normal(0, 4000, size=5003).astype(np.float32) for _ in range(3)]
    for x, y in zip(orc.mct(*c, inverse=True, real=True), refgrok.mct(*c, inverse=True, real=True)):
        assert np.array_equal(x.view(np.int32), y.view(np.int32))


@pytest.mark.parametrize("w,h,scale", [(64, 64, 40), (64, 64, 3), (32, 32, 20000), (17, 9, 3000), (64, 3, 300), (5, 64, 300), (1, 1, 9), (33, 37, 150)])
@pytest.mark.parametrize("orient", [0, 1, 2, 3])
def test_t1_decoder_against_the_reference_decoder(w, h, scale, orient):
    """orc_t1_decode_cblk against opj_t1_decode_cblk (t1.cpp:1337-1448) at every truncation point of the block"""
    rng = np.random.default_rng(w * 977 + h * 13 + scale + orient)
    c = rng.laplace(0, scale, size=(h, w)).astype(np.int32)
    if scale <= 5:
        c[rng.random((h, w)) < 0.6] = 0
    q, mx = orc.quantize(c)
    r = orc.t1_encode(q, orient, mx)
    for npasses in range(1, r["npasses"] + 1):
        data = r["data"][:r["rates"][npasses - 1]]
        if not data:
            continue
        got = orc.t1_decode(data, w, h, orient, r["numbps"], npasses)
        ref = refgrok.t1_decode_cblk(data, w, h, orient, r["numbps"], [(len(data), npasses)])
        assert np.array_equal(got, ref), "npasses %d of %d" % (npasses, r["npasses"])


@pytest.mark.parametrize("shape,numres", [((64, 64), 6), ((37, 53), 4), ((130, 257), 6), ((200, 31), 3)])
def test_dwt97_inverse_float_tight(shape, numres):
    """the float 9/7 inverse of the oracle against opj_dwt_decode_real: the same operations in the same order, so the
    results agree to the bit (the round-1 test allowed 2e-3 of the range)"""
    a = np.random.default_rng(3).normal(0, 500, size=shape).astype(np.float32)
    got = orc.dwt(a, numres, inverse=True, real=True)
    ref = refgrok.dwt_decode(a.view(np.int32), numres, real=1).view(np.float32)
    ulp = np.abs(got.view(np.int32).astype(np.int64) - ref.view(np.int32).astype(np.int64))
    assert ulp.max() <= 0, "max difference %d ulp" % ulp.max()
