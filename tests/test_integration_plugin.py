"""The drop-in boundary EXECUTED: the reference's unchanged plug-in layer (src/common: Codec registry, CopyBuffer,
RGBAoutputFile / RGBAinputFile; ext/Little-CMS) compiled for Linux into oracle/_ref/libj2kplugin.so by oracle/Makefile,
with j2k::CudaCodec (integration/j2k_cuda_codec.cpp over the C ABI) as the codec and POSIX files underneath
(integration/posix_io.cpp).  CPU part: the reference's own Codec::CopyBuffer (src/common/j2k_codec.cpp:222-427) is the oracle
of the sample conversion the MCT kernels run (j2k_b200/csrc/copybuffer.cuh, compiled for the host by tests/native).
GPU part: After Effects style ARGB worlds written and read back through the reference's RGBA layer."""
import ctypes as C
import os
import subprocess
import numpy as np
import pytest
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame

HERE = os.path.dirname(os.path.abspath(__file__))
PLUGIN = os.path.join(HERE, "..", "oracle", "_ref", "libj2kplugin.so")
needs_plugin = pytest.mark.skipif(not os.path.exists(PLUGIN), reason="oracle/_ref/libj2kplugin.so not built (reference sources absent)")
TYPES = {0: np.uint8, 1: np.uint16, 2: np.uint32, 3: np.int32}
LOSSLESS, SIZE, QUALITY, CINEMA = range(4)
J2C, JP2 = 1, 2


@pytest.fixture(scope="module")
def plug():
    lib = C.CDLL(PLUGIN)
    lib.itest_default_codec_name.restype = C.c_char_p
    return lib


@pytest.fixture(scope="module")
def native():
    subprocess.check_call(["make", "-C", os.path.join(HERE, "native")], stdout=subprocess.DEVNULL)
    return C.CDLL(os.path.join(HERE, "native", "libt1dec_host.so"))


def ref_copy(plug, src, st, sd, ss, dt, dd, ds):
    """the reference's Codec::CopyBuffer on a 2-D array"""
    src = np.ascontiguousarray(src, dtype=TYPES[st])
    dst = np.zeros(src.shape, dtype=TYPES[dt])
    h, w = src.shape
    plug.itest_copy_channel(dst.ctypes.data_as(C.c_void_p), dt, dd, ds, C.c_long(dst.itemsize), C.c_long(dst.strides[0]),
                            src.ctypes.data_as(C.c_void_p), st, sd, ss, C.c_long(src.itemsize), C.c_long(src.strides[0]), w, h)
    return dst


CONVS = [  # src type, src depth, src signed, dst type, dst depth, dst signed
    (0, 8, 0, 3, 8, 0), (1, 16, 0, 3, 16, 0), (1, 16, 0, 3, 12, 0), (0, 8, 0, 3, 12, 0), (1, 12, 0, 3, 16, 0), (1, 10, 0, 3, 8, 0),
    (0, 8, 0, 3, 16, 0), (0, 8, 0, 3, 15, 0), (1, 9, 0, 3, 16, 0),
    (3, 12, 0, 1, 16, 0), (3, 16, 0, 1, 16, 0), (3, 8, 0, 0, 8, 0), (3, 12, 0, 0, 8, 0), (3, 10, 0, 1, 16, 0), (3, 8, 0, 1, 16, 0),
    (3, 8, 0, 2, 24, 0), (3, 8, 0, 2, 17, 0), (3, 9, 0, 3, 31, 0), (3, 8, 0, 3, 20, 0),             # two fills (j2k_codec.cpp:302-325)
    (3, 1, 0, 0, 8, 0), (3, 2, 0, 0, 8, 0), (3, 4, 0, 1, 16, 0), (3, 5, 0, 1, 16, 0), (3, 7, 0, 0, 8, 0), (3, 3, 0, 1, 12, 0), (0, 4, 0, 3, 12, 0),  # depth < 8 (:326-355)
    (3, 8, 1, 0, 8, 0), (3, 16, 1, 1, 16, 0), (3, 12, 1, 1, 16, 0), (3, 12, 1, 0, 8, 0), (3, 5, 1, 0, 8, 0), (3, 12, 1, 3, 12, 1), (3, 12, 1, 3, 16, 1),  # signed components
    (2, 20, 0, 3, 16, 0), (2, 32, 0, 3, 16, 0), (3, 16, 0, 2, 32, 0),
]


@needs_plugin
@pytest.mark.parametrize("st,sd,ss,dt,dd,ds", CONVS)
def test_sample_conversion_equals_the_reference_copybuffer(plug, native, st, sd, ss, dt, dd, ds):
    rng = np.random.default_rng(sd * 100 + dd)
    lo, hi = (-(1 << (sd - 1)), (1 << (sd - 1))) if ss else (0, 1 << sd)
    vals = rng.integers(lo, hi, size=(7, 501), dtype=np.int64)
    vals[0, :4] = [lo, hi - 1, lo + 1 if hi - lo > 1 else lo, (lo + hi) // 2]
    want = ref_copy(plug, vals.astype(TYPES[st]), st, sd, ss, dt, dd, ds)
    src64 = vals.astype(TYPES[st]).astype(np.int64).ravel().copy()      # as loaded: zero-extended unless INT
    out64 = np.zeros_like(src64)
    native.copyconv_host(src64.ctypes.data_as(C.c_void_p), out64.ctypes.data_as(C.c_void_p), src64.size, st, sd, ss, dt, dd, ds)
    got = out64.astype(np.uint64).astype(np.uint32).view(np.int32).astype(TYPES[dt]) if dt == 3 else out64.astype(TYPES[dt])
    assert np.array_equal(got.reshape(want.shape), want)


@needs_plugin
def test_reference_codec_registry_links_and_runs(plug):
    assert plug.itest_default_codec_name() == b"OpenJPEG"           # the unchanged CodecContainer (j2k_codec.cpp:508-519)


# ---- GPU: worlds through RGBAoutputFile -> CudaCodec -> file -> RGBAinputFile ----
def _write(plug, path, world, channels, depth, fmt=J2C, method=LOSSLESS, layers=1, tile=0, size_kib=50, quality=50, reversible=1,
           dci4k=0, as_written=0):
    h, w, _ = world.shape
    err = C.create_string_buffer(512)
    rc = plug.itest_write(path.encode(), world.ctypes.data_as(C.c_void_p), w, h, world.strides[0], world.itemsize, channels, depth, fmt,
                          method, layers, tile, size_kib, quality, reversible, dci4k, as_written, err, 512)
    assert rc == 0, err.value.decode()


def _read(plug, path, shape, dtype, subsample=1):
    world = np.zeros(shape, dtype=dtype)
    info = (C.c_int * 8)()
    err = C.create_string_buffer(512)
    rc = plug.itest_read(path.encode(), world.ctypes.data_as(C.c_void_p), shape[1], shape[0], world.strides[0], world.itemsize, subsample,
                         info, err, 512)
    assert rc == 0, err.value.decode()
    return world, list(info)


def _world(h, w, dtype, seed, kind="smooth"):
    """A, R, G, B interleaved, as PF_EffectWorld (src/aftereffects/j2k.cpp:324-362)"""
    prec = 8 if dtype == np.uint8 else 16
    return np.ascontiguousarray(make_frame(h, w, 4, prec, kind, seed).astype(dtype))


@needs_plugin
@pytest.mark.gpu
@pytest.mark.parametrize("dtype,channels,fmt", [(np.uint8, 3, J2C), (np.uint8, 4, J2C), (np.uint16, 3, J2C), (np.uint16, 4, JP2), (np.uint8, 4, JP2)])
def test_argb_world_round_trip_through_the_reference_rgba_layer(plug, tmp_path, dtype, channels, fmt):
    world = _world(123, 200, dtype, 5)
    path = str(tmp_path / "f.j2k")
    _write(plug, path, world, channels, 8 * world.itemsize, fmt=fmt)
    back, info = _read(plug, path, world.shape, dtype)
    assert info[:4] == [200, 123, channels, 8 * world.itemsize] and info[4] == fmt
    assert np.array_equal(back[..., 1:], world[..., 1:])             # R, G, B bit-exact
    if channels == 4:
        assert np.array_equal(back[..., 0], world[..., 0])           # alpha carried as the fourth component
    else:
        # RGBAinputFile fills a missing alpha with "white" -- through a cast to uint8_t (j2k_rgba_file.cpp:416), so a 16-bit
        # world gets 255, not 65535: the reference's behaviour, unchanged above the boundary
        assert np.all(back[..., 0] == 255)


@needs_plugin
@pytest.mark.gpu
def test_subsampled_read_and_second_call_reuses_the_context(plug, tmp_path):
    import j2k
    world = _world(256, 384, np.uint16, 9)
    path = str(tmp_path / "f.j2k")
    for _ in range(2):                                               # the second WriteFile / ReadFile find their contexts in the cache
        _write(plug, path, world, 3, 16)
        half, _ = _read(plug, path, (128, 192, 4), np.uint16, subsample=2)
    full = j2k.decode(open(path, "rb").read(), reduce=1, dtype=np.uint16)
    assert np.array_equal(half[..., 1:], full)                       # = the C ABI's reduce = 1 (checked against the reference decoder elsewhere)


@needs_plugin
@pytest.mark.gpu
def test_cinema_write_of_a_16_bit_world_as_12_bit_codestream(plug, tmp_path):
    """the plug-in's cinema path: a 16-bit world, FileInfo.depth = 12 (src/aftereffects/j2k.cpp:843-847): CopyBuffer's down-shift"""
    import j2k
    world = _world(540, 1024, np.uint16, 2)
    path = str(tmp_path / "c.j2c")
    _write(plug, path, world, 3, 12, method=CINEMA, size_kib=1271, dci4k=0)
    data = open(path, "rb").read()
    info = j2k.probe(data)
    assert info.prec[0] == 12 and info.profile == L.PROFILE_CINEMA_2K and not info.reversible
    back, _ = _read(plug, path, world.shape, np.uint16)              # 12-bit components up-shifted with bit replication to 16
    want = world[..., 1:] >> 4
    got12 = back[..., 1:] >> 4
    err = np.abs(got12.astype(np.int32) - want.astype(np.int32))
    assert err.mean() < 2.0 and np.array_equal(back[..., 1:], (got12 << 4) | (got12 >> 8))


@needs_plugin
@pytest.mark.gpu
@pytest.mark.parametrize("tile", [0, 128])
def test_as_written_mode_is_byte_identical_to_the_reference_codec(plug, tmp_path, tile):
    """AS_WRITTEN = what OpenJPEGCodec::WriteFile passes on today (j2k_openjpeg_codec.cpp:703-719): 12 lossless layers, no MCT,
    tile size; the reference codec core (ext/grok compiled in oracle/_ref) with those parameters must write the same bytes"""
    import refgrok
    if not refgrok.available():
        pytest.skip("reference not built")
    world = _world(300, 420, np.uint8, 4)
    path = str(tmp_path / "w.j2c")
    _write(plug, path, world, 3, 8, layers=12, tile=tile, as_written=1)
    ours = open(path, "rb").read()
    ref = refgrok.encode([world[..., c].astype(np.int32) for c in (1, 2, 3)], 8, mct=0, numlayers=12, tile=(tile, tile))
    assert ours == ref


# ---- GPU: the C ABI's own conversion against the reference's CopyBuffer ----
@needs_plugin
@pytest.mark.gpu
@pytest.mark.parametrize("st,sd,prec", [(0, 8, 12), (1, 16, 12), (1, 12, 16), (1, 10, 8), (0, 4, 12), (2, 20, 16), (3, 12, 12)])
def test_encode_with_depth_other_than_precision(plug, st, sd, prec):
    """j2kcuda_encode(planes of depth sd, codestream precision prec) == encode of CopyBuffer(planes) at matching depth"""
    import j2k
    rng = np.random.default_rng(sd + prec)
    img = rng.integers(0, 1 << sd, size=(90, 131, 3), dtype=np.int64).astype(TYPES[st])
    conv = np.stack([ref_copy(plug, img[..., c], st, sd, 0, 3, prec, 0) for c in range(3)], axis=-1)
    assert conv.min() >= 0 and conv.max() < (1 << prec)

    def enc(a, stype, depth):
        h, w, nc = a.shape
        p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, mct=1)
        pl = j2k.planes_of(np.ascontiguousarray(a))
        for c in range(nc):
            pl[c].sample_type = stype; pl[c].depth = depth
        out = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
        assert L.lib().j2kcuda_encode(C.byref(p), pl, nc, C.byref(out), C.byref(n)) == 0, L.last_error()
        data = C.string_at(out, n.value); L.lib().j2kcuda_free(out)
        return data
    a = np.ascontiguousarray(img)
    assert enc(a, st, sd) == enc(np.ascontiguousarray(conv.astype(np.int32)), 3, prec)


@needs_plugin
@pytest.mark.gpu
@pytest.mark.parametrize("prec,sgnd,dt,dd,ds", [(12, 0, 1, 16, 0), (8, 0, 1, 16, 0), (10, 0, 0, 8, 0), (5, 0, 0, 8, 0), (3, 0, 1, 12, 0), (12, 1, 1, 16, 0),
                                               (8, 1, 0, 8, 0), (16, 0, 2, 32, 0), (8, 0, 2, 24, 0), (12, 0, 3, 12, 0), (12, 1, 3, 12, 1)])
def test_decode_into_planes_of_another_depth_type_and_sign(plug, prec, sgnd, dt, dd, ds):
    """j2kcuda_decode into caller planes == the reference's CopyBuffer applied to the decoded int components
    (OpenJPEGCodec::ReadFile, j2k_openjpeg_codec.cpp:528-562)"""
    import j2k
    rng = np.random.default_rng(prec * 7 + dd)
    lo, hi = (-(1 << (prec - 1)), 1 << (prec - 1)) if sgnd else (0, 1 << prec)
    comps = rng.integers(lo, hi, size=(70, 99, 3), dtype=np.int64).astype(np.int32)
    h, w, nc = comps.shape
    p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, sgnd=sgnd, mct=0)
    pl = j2k.planes_of(np.ascontiguousarray(comps))
    for c in range(nc):
        pl[c].sample_type = L.INT; pl[c].depth = prec; pl[c].sgnd = sgnd
    out = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
    assert L.lib().j2kcuda_encode(C.byref(p), pl, nc, C.byref(out), C.byref(n)) == 0, L.last_error()
    cs = C.string_at(out, n.value); L.lib().j2kcuda_free(out)
    dst = np.zeros((h, w, nc), dtype=TYPES[dt])
    dp = j2k.planes_of(dst)
    for c in range(nc):
        dp[c].sample_type = dt; dp[c].depth = dd; dp[c].sgnd = ds
    assert L.lib().j2kcuda_decode(cs, C.c_size_t(len(cs)), 0, dp, nc, None, None) == 0, L.last_error()
    for c in range(nc):
        assert np.array_equal(dst[..., c], ref_copy(plug, comps[..., c], 3, prec, sgnd, dt, dd, ds)), c
