"""GPU parity, stage by stage through the C ABI: CUDA kernels vs the oracle on the same seeded inputs.
Integer stages are bit-exact."""
import numpy as np
import pytest
import orc
import j2k

pytestmark = pytest.mark.gpu


def rnd(shape, lo, hi, seed):
    return np.random.default_rng(seed).integers(lo, hi, size=shape, dtype=np.int64).astype(np.int32)


def first_diff(a, b):
    d = np.argwhere(a != b)
    return "%d diffs, first at %s: got %s want %s" % (len(d), d[0], a[tuple(d[0])], b[tuple(d[0])]) if len(d) else "equal"


@pytest.mark.parametrize("irreversible", [0, 1])
def test_mct_forward(irreversible):
    c = [rnd((100003,), -(1 << 17), 1 << 17, s) * (2048 if irreversible else 1) for s in range(3)]
    got = j2k.stage_mct(*c, irreversible=irreversible)
    want = orc.mct(*c, real=bool(irreversible))
    for g, w in zip(got, want):
        assert np.array_equal(g, w), first_diff(g, w)


def test_rct_inverse():
    c = [rnd((100003,), -(1 << 17), 1 << 17, s) for s in range(3)]
    for g, w in zip(j2k.stage_mct(*c, inverse=1), orc.mct(*c, inverse=True)):
        assert np.array_equal(g, w), first_diff(g, w)


def test_ict_inverse_float():
    c = [np.random.default_rng(s).normal(0, 3000, 50001).astype(np.float32) for s in range(3)]
    got = j2k.stage_mct(*[x.view(np.int32) for x in c], irreversible=1, inverse=1)
    want = orc.mct(*c, inverse=True, real=True)
    for g, w in zip(got, want):
        assert np.array_equal(g.view(np.float32), w)


SHAPES = [((64, 64), 0, 0, 6), ((37, 53), 0, 0, 4), ((33, 17), 3, 5, 3), ((1, 40), 0, 0, 3), ((40, 1), 1, 1, 3),
          ((130, 257), 0, 0, 6), ((2, 2), 0, 0, 2), ((3, 500), 1, 0, 3), ((270, 481), 0, 0, 6), ((256, 1024), 1024, 2048, 6),
          ((97, 131), 7, 9, 6)]


@pytest.mark.parametrize("shape,x0,y0,numres", SHAPES)
@pytest.mark.parametrize("irreversible", [0, 1])
def test_dwt_forward(shape, x0, y0, numres, irreversible):
    a = rnd(shape, -(1 << 16), 1 << 16, 11) * (256 if irreversible else 1)
    got = j2k.stage_dwt(a, numres, irreversible=irreversible, x0=x0, y0=y0)
    want = orc.dwt(a, numres, real=bool(irreversible), x0=x0, y0=y0)
    assert np.array_equal(got, want), first_diff(got, want)


@pytest.mark.parametrize("shape,x0,y0,numres", SHAPES)
def test_dwt53_inverse(shape, x0, y0, numres):
    a = rnd(shape, -(1 << 16), 1 << 16, 13)
    got = j2k.stage_dwt(a, numres, inverse=1, x0=x0, y0=y0)
    want = orc.dwt(a, numres, inverse=True, x0=x0, y0=y0)
    assert np.array_equal(got, want), first_diff(got, want)


@pytest.mark.parametrize("shape,x0,y0,numres", SHAPES)
def test_dwt97_inverse_float(shape, x0, y0, numres):
    a = np.random.default_rng(17).normal(0, 800, size=shape).astype(np.float32)
    got = j2k.stage_dwt(a.view(np.int32), numres, irreversible=1, inverse=1, x0=x0, y0=y0).view(np.float32)
    want = orc.dwt(a, numres, inverse=True, real=True, x0=x0, y0=y0)
    assert np.array_equal(got, want), "max abs diff %g" % float(np.abs(got - want).max())


def test_dwt53_roundtrip_4k():
    a = rnd((2160, 4096), -(1 << 16), 1 << 16, 19)
    f = j2k.stage_dwt(a, 6)
    assert np.array_equal(j2k.stage_dwt(f, 6, inverse=1), a)


def _check_blocks(plane, blocks, irreversible=0):
    got = j2k.stage_t1_encode(plane, blocks, irreversible)
    for (x, y, w, h, orient, bc), g in zip(blocks, got):
        q, mx = orc.quantize(plane[y:y + h, x:x + w], qmfbid=0 if irreversible else 1, bandconst=bc)
        want = orc.t1_encode(q, orient, mx)
        tag = "block (%d,%d %dx%d orient %d)" % (x, y, w, h, orient)
        assert g["numbps"] == want["numbps"], tag
        assert g["npasses"] == want["npasses"], tag
        assert g["nmsedec"] == want["nmsedec"], tag + " nmsedec %s vs %s" % (g["nmsedec"][:6], want["nmsedec"][:6])
        assert g["rates"] == want["rates"], tag + " rates %s vs %s" % (g["rates"][:8], want["rates"][:8])
        assert g["data"] == want["data"], tag


@pytest.mark.parametrize("scale", [2, 40, 3000, 60000])
def test_t1_blocks_64(scale):
    rng = np.random.default_rng(scale)
    plane = rng.laplace(0, scale, size=(128, 256)).astype(np.int32)
    if scale == 2:
        plane[rng.random(plane.shape) < 0.8] = 0
    blocks = [(x, y, 64, 64, (x // 64 + y // 64) % 4, 0) for y in (0, 64) for x in (0, 64, 128, 192)]
    _check_blocks(plane, blocks)


def test_t1_ragged_blocks():
    rng = np.random.default_rng(4)
    plane = rng.laplace(0, 500, size=(70, 200)).astype(np.int32)
    blocks = [(0, 0, 32, 32, 0, 0), (32, 0, 17, 9, 1, 0), (60, 3, 64, 3, 2, 0), (130, 0, 5, 64, 3, 0), (140, 0, 1, 1, 0, 0),
              (150, 0, 33, 37, 1, 0), (0, 40, 64, 30, 2, 0), (190, 60, 10, 10, 3, 0), (100, 40, 2, 7, 0, 0)]
    _check_blocks(plane, blocks)


def test_t1_zero_and_smooth_blocks():
    plane = np.zeros((64, 192), np.int32)
    y, x = np.mgrid[0:64, 0:64]
    plane[:, 64:128] = (x * 3 + y * 5 - 250)
    plane[:, 128:192] = ((x // 8 + y // 8) % 2) * 1000 - 500
    _check_blocks(plane, [(0, 0, 64, 64, 0, 0), (64, 0, 64, 64, 1, 0), (128, 0, 64, 64, 3, 0)])


def test_t1_irreversible_quantiser():
    rng = np.random.default_rng(8)
    plane = (rng.laplace(0, 300, size=(64, 128)) * 2048).astype(np.int32)
    _check_blocks(plane, [(0, 0, 64, 64, 1, 8192 * 8192 // 4000), (64, 0, 32, 32, 3, 8192 * 8192 // 16000)], irreversible=1)
