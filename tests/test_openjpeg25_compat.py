"""Interoperability with OpenJPEG 2.5.x, the only OpenJPEG that can run here (libopenjp2 2.5.4 inside Pillow, 2.5.3 inside
OpenCV): `ext/openjpeg`, the fork the plug-in links, is an empty directory in the reference snapshot.

  CPU   J2KCUDA_COMPAT_OPENJPEG_2_5 (every pass in a lossless layer + OpenJPEG's COM text) makes the host Tier-2 write the
        codestream OpenJPEG 2.5.4 writes, byte for byte (stages from the oracle, as in tests/test_t2_host.py)
  GPU   the same through the CUDA path; OpenJPEG-written codestreams / JP2 files decode bit-exact with libj2kcuda;
        libj2kcuda's streams decode bit-exact with OpenJPEG (8-bit through Pillow, 16-bit through OpenCV)."""
import io
import numpy as np
import pytest
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame
import hostpipe

Image = pytest.importorskip("PIL.Image")

CASES = [(96, 128, 3, "smooth", 6, 64), (130, 200, 3, "noise", 6, 64), (200, 333, 1, "smooth", 6, 64), (64, 64, 4, "smooth", 6, 64),
         (77, 301, 3, "ramp", 4, 32), (150, 90, 1, "noise", 3, 16), (33, 47, 4, "noise", 2, 64)]


def pillow_encode(img, numres, cblk, mct):
    nc = img.shape[2]
    buf = io.BytesIO()
    Image.fromarray(img.squeeze() if nc == 1 else img, {1: "L", 3: "RGB", 4: "RGBA"}[nc]).save(
        buf, "JPEG2000", no_jp2=True, irreversible=False, num_resolutions=numres, codeblock_size=(cblk, cblk), mct=mct, progression="LRCP")
    return buf.getvalue()


def pillow_decode(data):
    a = np.array(Image.open(io.BytesIO(data)))
    return a[..., None] if a.ndim == 2 else a


@pytest.mark.parametrize("h,w,nc,kind,numres,cblk", CASES)
def test_host_tier2_writes_openjpeg_2_5_bytes(h, w, nc, kind, numres, cblk):
    img = make_frame(h, w, nc, 8, kind, seed=7).astype(np.uint8)
    mct = 1 if nc == 3 else 0                       # Pillow turns the MCT on for three components only
    ref = pillow_encode(img, numres, cblk, mct)
    p = L.default_params(width=w, height=h, ncomp=nc, prec=8, mct=mct, numresolution=numres, cblk_w=cblk, cblk_h=cblk,
                         compat=L.COMPAT_OPENJPEG)
    ours = hostpipe.encode([img[..., c].astype(np.int32) for c in range(nc)], 8, p)
    assert ours == ref


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,nc,kind,numres,cblk", CASES + [(1080, 2048, 3, "smooth", 6, 64)])
def test_cuda_path_and_openjpeg_2_5_agree(h, w, nc, kind, numres, cblk):
    import j2k
    img = make_frame(h, w, nc, 8, kind, seed=7).astype(np.uint8)
    mct = 1 if nc == 3 else 0
    ref = pillow_encode(img, numres, cblk, mct)
    ours = j2k.encode(img, 8, mct=mct, numresolution=numres, cblk_w=cblk, cblk_h=cblk, compat=L.COMPAT_OPENJPEG)
    assert ours == ref                                               # byte-identical to OpenJPEG 2.5.4
    assert np.array_equal(j2k.decode(ref), img)                      # OpenJPEG's stream, our decoder
    grok_mode = j2k.encode(img, 8, mct=mct, numresolution=numres, cblk_w=cblk, cblk_h=cblk)
    assert np.array_equal(pillow_decode(grok_mode), img)             # our default stream, OpenJPEG's decoder


@pytest.mark.gpu
@pytest.mark.parametrize("nc", [3, 4])
def test_16_bit_through_opencv(nc):
    cv2 = pytest.importorskip("cv2")
    import j2k
    img = make_frame(270, 480, nc, 16, "smooth", seed=3)
    ok, enc = cv2.imencode(".jp2", img, [cv2.IMWRITE_JPEG2000_COMPRESSION_X1000, 1000])    # OpenJPEG 2.5.3, lossless, JP2 wrapper
    assert ok
    order = [2, 1, 0, 3][:nc]                                        # OpenCV's arrays are B, G, R(, A); the file's components R, G, B(, A)
    assert np.array_equal(j2k.decode(enc.tobytes())[..., order], img)
    ours = j2k.encode(img, 16, mct=1 if nc == 3 else 0)
    back = cv2.imdecode(np.frombuffer(ours, np.uint8), cv2.IMREAD_UNCHANGED)
    assert back is not None and back.dtype == np.uint16
    assert np.array_equal(back[..., order], img)


# ---- 16-bit byte identity with a real OpenJPEG (row N1): OpenCV's embedded libopenjp2 2.5.3 writes lossless 5/3, 6 resolutions,
# 64x64 code-blocks, LRCP, one layer, NO multi-component transform, inside a JP2 wrapper; its jp2c payload is the codestream ----
def opencv_codestream(img):
    cv2 = pytest.importorskip("cv2")
    import ctypes as C
    ok, enc = cv2.imencode(".jp2", img, [cv2.IMWRITE_JPEG2000_COMPRESSION_X1000, 1000])
    assert ok
    b = enc.tobytes()
    info = L.Info()
    assert L.lib().j2kcuda_probe(b, C.c_size_t(len(b)), C.byref(info)) == 0, L.last_error()      # host-only call
    assert info.is_jp2 and info.mct == 0 and info.numresolution == 6 and info.numlayers == 1 and info.prec[0] == 16
    return b[info.codestream_offset:info.codestream_offset + info.codestream_length]


@pytest.mark.parametrize("h,w,nc,kind", [(270, 480, 3, "smooth"), (270, 480, 4, "smooth"), (135, 211, 3, "noise"), (97, 64, 4, "ramp")])
def test_host_tier2_writes_openjpeg_2_5_3_bytes_for_16_bit(h, w, nc, kind):
    img = make_frame(h, w, nc, 16, kind, seed=3)
    ref = opencv_codestream(img)
    order = [2, 1, 0, 3][:nc]                                        # OpenCV's arrays are B, G, R(, A); the file's components R, G, B(, A)
    p = L.default_params(width=w, height=h, ncomp=nc, prec=16, mct=0, compat=L.COMPAT_OPENJPEG_253)
    assert hostpipe.encode([img[..., c].astype(np.int32) for c in order], 16, p) == ref


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,nc", [(270, 480, 3), (270, 480, 4), (2160, 4096, 3), (2160, 4096, 4)])
def test_cuda_16_bit_codestream_is_byte_identical_to_openjpeg_2_5_3(h, w, nc):
    """north_star: "4K 16-bit RGB lossless encode ... byte-exact with OpenJPEG" -- against the OpenJPEG that can run here"""
    import j2k
    img = make_frame(h, w, nc, 16, "smooth", seed=11)
    ref = opencv_codestream(img)
    order = [2, 1, 0, 3][:nc]
    ours = j2k.encode(np.ascontiguousarray(img[..., order]), 16, mct=0, compat=L.COMPAT_OPENJPEG_253)
    assert ours == ref
    assert np.array_equal(j2k.decode(ref)[..., order], img)          # and its stream through our decoder, bit-exact
