"""Generates tests/golden/*.npz from the REFERENCE ITSELF (ext/grok compiled into oracle/_ref/libgrokref.so by oracle/Makefile)
and from OpenJPEG 2.5.x (Pillow / OpenCV).  Run in the build container (needs /root/reference); the fixtures are committed so
that the parity tests hold fixed known answers wherever they run.   python tests/golden/make_golden.py"""
import io
import os
import sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
import refgrok
from j2k_b200.frames import make_frame


def u8(b):
    return np.frombuffer(b, dtype=np.uint8)


def main():
    rng = np.random.default_rng(2026)
    # ---- stage vectors: MCT, DWT, Tier-1 blocks ----
    c = [rng.integers(-3000, 3000, size=257).astype(np.int32) for _ in range(3)]
    cf = [rng.normal(0, 900, size=257).astype(np.float32) for _ in range(3)]
    plane = rng.integers(-2000, 2000, size=(45, 67)).astype(np.int32)
    planef = rng.normal(0, 300, size=(45, 67)).astype(np.float32)
    st = dict(mct_in=np.stack(c), rct_fwd=np.stack(refgrok.mct(*c)), rct_inv=np.stack(refgrok.mct(*c, inverse=True)),
              ict_fwd=np.stack(refgrok.mct(*[x * 2048 for x in c], real=True)), ict_inv_in=np.stack(cf),
              ict_inv=np.stack(refgrok.mct(*cf, inverse=True, real=True)),
              dwt_in=plane, dwt53_fwd=refgrok.dwt_encode(plane, 4, x0=3, y0=5), dwt53_inv=refgrok.dwt_decode(plane, 4, x0=3, y0=5),
              dwt97_fwd=refgrok.dwt_encode(plane * 8, 4, real=1), dwt97_inv_in=planef,
              dwt97_inv=refgrok.dwt_decode(planef.view(np.int32), 4, real=1).view(np.float32))
    blocks = []
    for i, (w, h, scale, orient) in enumerate([(64, 64, 40, 0), (32, 32, 900, 1), (17, 9, 3000, 2), (64, 3, 5, 3), (5, 64, 300, 1)]):
        b = rng.laplace(0, scale, size=(h, w)).astype(np.int32)
        r = refgrok.t1_encode_cblk(b, orient)
        full = refgrok.t1_decode_cblk(r["data"], w, h, orient, r["numbps"], [(len(r["data"]), r["npasses"])])
        cut = max(1, r["npasses"] // 2)
        part = refgrok.t1_decode_cblk(r["data"][:r["rates"][cut - 1]], w, h, orient, r["numbps"], [(r["rates"][cut - 1], cut)])
        st.update({"blk%d_in" % i: b, "blk%d_meta" % i: np.array([orient, r["numbps"], r["npasses"], cut], np.int32),
                   "blk%d_rates" % i: np.array(r["rates"], np.int32), "blk%d_dist" % i: np.array(r["dist"], np.float64),
                   "blk%d_bytes" % i: u8(r["data"]), "blk%d_dec" % i: full, "blk%d_dec_cut" % i: part})
    np.savez_compressed(os.path.join(HERE, "stages.npz"), **st)
    # ---- whole codestreams written by the reference encoder, and the pixels its decoder gives back ----
    cs = {}
    cases = [("c1_rgb8", 72, 136, 3, 8, "smooth", dict(mct=1)),
             ("c2_rgba16", 70, 130, 4, 16, "smooth", dict(mct=1)),
             ("c3_cinema2k", 1080, 2048, 3, 12, "smooth", dict(irreversible=1, mct=1, rsiz=3, max_cs_size=30000, max_comp_size=25000, numresolution=6, cblk=(32, 32), prog=4)),
             ("c4_tiles", 100, 150, 3, 16, "noise", dict(mct=1, tile=(64, 64), numresolution=4)),
             ("plugin_as_written", 90, 120, 3, 8, "ramp", dict(mct=0, numlayers=12, tile=(64, 64), numresolution=4)),
             ("layers_rate", 128, 128, 1, 8, "noise", dict(irreversible=1, numlayers=3, rates=(20, 8, 3)))]
    for name, h, w, nc, prec, kind, kw in cases:
        if name == "c3_cinema2k":
            img = make_frame(h, w, nc, prec, kind, 5)
        else:
            img = make_frame(h, w, nc, prec, kind, 5)
        planes = [img[..., k].astype(np.int32) for k in range(nc)]
        data = refgrok.encode(planes, prec, **kw)
        dec = np.stack(refgrok.decode(data), axis=-1)
        if name == "c3_cinema2k":            # keep the fixture small: store the 35 KB stream and a checksum of the pixels, not 13 MB of them
            cs[name + "_stream"] = u8(data)
            cs[name + "_pixsum"] = np.array([int(dec.astype(np.int64).sum()), int((dec.astype(np.int64) ** 2).sum() % (1 << 61))], np.int64)
            cs[name + "_row0"] = dec[0].astype(np.uint16)
        else:
            cs[name + "_img"] = img
            cs[name + "_stream"] = u8(data)
            if kw.get("irreversible"):
                cs[name + "_decoded"] = dec.astype(img.dtype)       # lossy: what the reference decoder gives back
            else:
                assert np.array_equal(dec, img)                     # lossless: the reference decoder returns the input
    # ---- OpenJPEG 2.5.x streams (the only OpenJPEG that runs here) ----
    try:
        from PIL import Image
        img = make_frame(60, 100, 3, 8, "smooth", 8).astype(np.uint8)
        buf = io.BytesIO()
        Image.fromarray(img, "RGB").save(buf, "JPEG2000", no_jp2=True, irreversible=False, num_resolutions=6, codeblock_size=(64, 64), mct=1, progression="LRCP")
        cs["opj254_rgb8_img"] = img; cs["opj254_rgb8_stream"] = u8(buf.getvalue())
    except Exception as e:
        print("Pillow fixture skipped:", e)
    try:
        import cv2
        img = make_frame(60, 100, 3, 16, "smooth", 9)
        ok, enc = cv2.imencode(".jp2", img, [cv2.IMWRITE_JPEG2000_COMPRESSION_X1000, 1000])
        cs["opj253_rgb16_img"] = img; cs["opj253_rgb16_jp2"] = enc.copy()
    except Exception as e:
        print("OpenCV fixture skipped:", e)
    np.savez_compressed(os.path.join(HERE, "codestreams.npz"), **cs)
    for f in ("stages.npz", "codestreams.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
