"""ctypes binding of libj2kcuda.so (include/j2kcuda.h).  Fails loudly when the library is missing:
there is no Python or CPU fallback for the pixel path."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("J2KCUDA_LIB", os.path.join(_HERE, "libj2kcuda.so"))

OK, ERR_PARAM, ERR_NO_DEVICE, ERR_CUDA, ERR_STREAM, ERR_NOMEM, ERR_ABORTED, ERR_UNSUPPORTED = range(8)
UCHAR, USHORT, UINT, INT = range(4)
LRCP, RLCP, RPCL, PCRL, CPRL = range(5)
PROFILE_NONE, PROFILE_CINEMA_2K, PROFILE_CINEMA_4K = 0, 3, 4
COMPAT_GROK, COMPAT_OPENJPEG, COMPAT_OPENJPEG_253 = 0, 1, 2


class Plane(C.Structure):
    _fields_ = [("buf", C.c_void_p), ("colbytes", C.c_int64), ("rowbytes", C.c_int64),
                ("width", C.c_uint32), ("height", C.c_uint32), ("sample_type", C.c_uint8),
                ("depth", C.c_uint8), ("sgnd", C.c_uint8), ("reserved", C.c_uint8)]


class Params(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("ncomp", C.c_uint32),
                ("prec", C.c_uint32), ("sgnd", C.c_uint32), ("irreversible", C.c_uint32),
                ("numresolution", C.c_uint32), ("cblk_w", C.c_uint32), ("cblk_h", C.c_uint32),
                ("mct", C.c_uint32), ("prog_order", C.c_uint32), ("tile_w", C.c_uint32),
                ("tile_h", C.c_uint32), ("numlayers", C.c_uint32), ("profile", C.c_uint32),
                ("max_comp_size", C.c_uint32), ("max_cs_size", C.c_uint64),
                ("rates", C.c_double * 16), ("compat", C.c_uint32), ("device", C.c_int32),
                ("tile_first", C.c_uint32), ("tile_count", C.c_uint32),
                ("cblksty", C.c_uint32), ("reserved", C.c_uint32 * 3)]


class Info(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("ncomp", C.c_uint32),
                ("prec", C.c_uint32 * 4), ("sgnd", C.c_uint32 * 4), ("dx", C.c_uint32 * 4),
                ("dy", C.c_uint32 * 4), ("reversible", C.c_uint32), ("numresolution", C.c_uint32),
                ("numlayers", C.c_uint32), ("mct", C.c_uint32), ("prog_order", C.c_uint32),
                ("tile_w", C.c_uint32), ("tile_h", C.c_uint32), ("tiles_x", C.c_uint32),
                ("tiles_y", C.c_uint32), ("cblk_w", C.c_uint32), ("cblk_h", C.c_uint32),
                ("profile", C.c_uint32), ("is_jp2", C.c_uint32), ("codestream_offset", C.c_uint64),
                ("codestream_length", C.c_uint64)]


JP2_MAX_CHANNELS = 16


class Jp2Info(C.Structure):
    _fields_ = [("is_jp2", C.c_uint32), ("brand", C.c_uint32), ("minversion", C.c_uint32),
                ("width", C.c_uint32), ("height", C.c_uint32), ("ncomp", C.c_uint32),
                ("bpc", C.c_uint32), ("unk_c", C.c_uint32), ("ipr", C.c_uint32),
                ("bpcc", C.c_uint32 * JP2_MAX_CHANNELS), ("colr_method", C.c_uint32),
                ("enumcs", C.c_uint32), ("precedence", C.c_uint32), ("approx", C.c_uint32),
                ("icc_offset", C.c_uint64), ("icc_length", C.c_uint64), ("cdef_n", C.c_uint32),
                ("cdef_cn", C.c_uint16 * JP2_MAX_CHANNELS), ("cdef_typ", C.c_uint16 * JP2_MAX_CHANNELS),
                ("cdef_asoc", C.c_uint16 * JP2_MAX_CHANNELS), ("pclr_entries", C.c_uint32),
                ("pclr_channels", C.c_uint32), ("pclr_depth", C.c_uint8 * JP2_MAX_CHANNELS),
                ("pclr_offset", C.c_uint64), ("cmap_n", C.c_uint32),
                ("cmap_cmp", C.c_uint16 * JP2_MAX_CHANNELS), ("cmap_mtyp", C.c_uint8 * JP2_MAX_CHANNELS),
                ("cmap_pcol", C.c_uint8 * JP2_MAX_CHANNELS), ("has_capture_res", C.c_uint32),
                ("has_display_res", C.c_uint32), ("capture_res", C.c_double * 2),
                ("display_res", C.c_double * 2), ("codestream_offset", C.c_uint64),
                ("codestream_length", C.c_uint64)]


class Jp2Params(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("ncomp", C.c_uint32),
                ("prec", C.c_uint32 * JP2_MAX_CHANNELS), ("sgnd", C.c_uint32 * JP2_MAX_CHANNELS),
                ("enumcs", C.c_uint32), ("icc", C.c_char_p), ("icc_len", C.c_size_t),
                ("alpha_comp", C.c_int32)]


class BlockResult(C.Structure):
    _fields_ = [("data", C.POINTER(C.c_uint8)), ("numbps", C.c_int32), ("npasses", C.c_int32),
                ("rates", C.POINTER(C.c_int32)), ("dist", C.POINTER(C.c_double))]


_lib = None


def lib():
    """Load libj2kcuda.so; raise if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libj2kcuda.so not built at %s: run `make -C j2k_b200/csrc` "
                               "(there is no fallback path)" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        l.j2kcuda_last_error.restype = C.c_char_p
        l.j2kcuda_free.argtypes = [C.c_void_p]
        l.j2kcuda_free.restype = None
        for name in ("j2kcuda_encoder_create", "j2kcuda_decoder_create"):
            if hasattr(l, name):
                getattr(l, name).restype = C.c_void_p
        _lib = l
    return _lib


def last_error():
    return (lib().j2kcuda_last_error() or b"").decode("utf-8", "replace")


def default_params(**kw):
    p = Params()
    lib().j2kcuda_default_params(C.byref(p))
    for k, v in kw.items():
        if k == "rates":
            for i, r in enumerate(v):
                p.rates[i] = r
        else:
            setattr(p, k, v)
    return p
