"""j2kcuda_dwt_forward_noalloc: the forward DWT for callers that cannot call cudaMalloc (include/j2kcuda.h; no counterpart in the
reference).  Its ping-pong plane and job table live in a fixed device array inside
the library.  Checked through the C ABI against the existing entry point j2kcuda_stage_dwt (bit-exact: 5/3 and the fixed-point
9/7 are integer arithmetic), on calls in a row that share the scratch array, with different shapes so that stale contents of
the previous call would show.  Collected last (test_zz_*)."""
import ctypes as C
import numpy as np
import pytest
from j2k_b200 import _lib as L

pytestmark = pytest.mark.gpu
MAX_SCRATCH = 2 * 1024 * 1024          # J2KCUDA_NOALLOC_MAX_SCRATCH_SAMPLES


def _proto(lib):
    lib.j2kcuda_dwt_forward_noalloc.restype = C.c_int
    lib.j2kcuda_dwt_forward_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.j2kcuda_stage_dwt.restype = C.c_int
    lib.j2kcuda_stage_dwt.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]


def _existing(lib, plane, x0, y0, numres, irreversible):
    """the existing entry point: host array in and out, Mallat layout"""
    ref = np.ascontiguousarray(plane, dtype=np.int32).copy()
    h, w = ref.shape
    assert lib.j2kcuda_stage_dwt(ref.ctypes.data, x0, y0, x0 + w, y0 + h, numres, irreversible, 0) == 0, L.last_error()
    return ref


def _noalloc(lib, torch, plane, x0, y0, numres, irreversible, extra_pitch=0, warm=False):
    """the new entry point: device planes allocated HERE by the caller (torch), none inside the call"""
    h, w = plane.shape
    pitch = ((w + 3) & ~3) + extra_pitch
    src = torch.full((h, pitch), 0x5a5a5a5, dtype=torch.int32, device="cuda")
    src[:, :w] = torch.from_numpy(np.ascontiguousarray(plane, dtype=np.int32)).cuda()
    keep = src.clone()
    dst = torch.full((h, pitch), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.j2kcuda_dwt_forward_noalloc(src.data_ptr(), dst.data_ptr(), pitch, x0, y0, x0 + w, y0 + h, numres, irreversible, None)
    assert rc == 0, L.last_error()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    if warm:      # once the module is resident (lazy loading maps kernels and the static array on first use): not a byte less free memory
        assert torch.cuda.mem_get_info()[0] >= free0
    assert torch.equal(src, keep)                                  # the input plane is not modified
    return dst[:, :w].cpu().numpy()


CASES = [  # (h, w, x0, y0, numres, irreversible)
    (1080, 2048, 0, 0, 6, 0),      # BASELINE C1's component size; 5 levels: the last LL is written straight into the output
    (270, 517, 0, 0, 5, 0),        # 4 levels: the last LL comes back from the scratch plane
    (1024, 1024, 0, 0, 6, 1),      # 9/7 fixed point
    (333, 257, 3, 5, 4, 0),        # odd tile origin: the parity of the origin decides which samples are low-pass
    (64, 80, 1, 1, 7, 1),          # deeper than the plane: the last levels are a few samples
    (37, 41, 0, 0, 2, 0),          # one level: the scratch plane is not used
    (16, 16, 0, 0, 1, 0),          # no level: a copy
]


def test_noalloc_forward_dwt_matches_the_existing_entry_point_on_calls_in_a_row():
    import torch
    lib = L.lib()
    _proto(lib)
    rng = np.random.default_rng(2024)
    # two rounds over every case, back to back in one process: every call after the first finds the previous call's
    # coefficients in the shared scratch array
    for rnd in range(2):
        for (h, w, x0, y0, numres, irr) in CASES:
            plane = rng.integers(-(1 << 15), 1 << 15, size=(h, w), dtype=np.int32)
            want = _existing(lib, plane, x0, y0, numres, irr)
            got = _noalloc(lib, torch, plane, x0, y0, numres, irr, extra_pitch=8 * rnd, warm=rnd == 1)
            assert np.array_equal(got, want), (rnd, h, w, x0, y0, numres, irr)


def test_noalloc_same_input_twice_gives_the_same_output():
    import torch
    lib = L.lib()
    _proto(lib)
    rng = np.random.default_rng(7)
    big = rng.integers(-4096, 4096, size=(2048, 2048), dtype=np.int32)          # exactly the stated maximum
    small = rng.integers(-4096, 4096, size=(100, 300), dtype=np.int32)
    a = _noalloc(lib, torch, big, 0, 0, 6, 0)
    s = _noalloc(lib, torch, small, 0, 0, 6, 1)                                   # dirties the scratch in between
    b = _noalloc(lib, torch, big, 0, 0, 6, 0)
    assert np.array_equal(a, b)
    assert np.array_equal(a, _existing(lib, big, 0, 0, 6, 0))
    assert np.array_equal(s, _existing(lib, small, 0, 0, 6, 1))


def test_noalloc_rejects_planes_beyond_the_fixed_scratch_and_bad_arguments():
    import torch
    lib = L.lib()
    _proto(lib)
    h, w = 2050, 2048                                                            # pitch * ceil(h / 2) = 2048 * 1025 > 2 Mi samples
    src = torch.zeros((h, w), dtype=torch.int32, device="cuda")
    dst = torch.zeros((h, w), dtype=torch.int32, device="cuda")
    rc = lib.j2kcuda_dwt_forward_noalloc(src.data_ptr(), dst.data_ptr(), w, 0, 0, w, h, 6, 0, None)
    assert rc == L.ERR_UNSUPPORTED
    msg = L.last_error()
    assert "J2KCUDA_NOALLOC_MAX_SCRATCH_SAMPLES" in msg and str(MAX_SCRATCH) in msg and str(2048 * 1025) in msg
    # one level needs no scratch plane: the same plane is fine
    assert lib.j2kcuda_dwt_forward_noalloc(src.data_ptr(), dst.data_ptr(), w, 0, 0, w, h, 2, 0, None) == 0, L.last_error()
    torch.cuda.synchronize()
    assert lib.j2kcuda_dwt_forward_noalloc(None, dst.data_ptr(), w, 0, 0, w, h, 2, 0, None) == L.ERR_PARAM
    assert lib.j2kcuda_dwt_forward_noalloc(src.data_ptr(), src.data_ptr(), w, 0, 0, w, h, 2, 0, None) == L.ERR_PARAM      # aliased
    assert lib.j2kcuda_dwt_forward_noalloc(src.data_ptr(), dst.data_ptr(), w - 2, 0, 0, w, h, 2, 0, None) == L.ERR_PARAM  # pitch < width
    assert lib.j2kcuda_dwt_forward_noalloc(src.data_ptr(), dst.data_ptr(), w + 2, 0, 0, w, h, 2, 0, None) == L.ERR_PARAM  # pitch % 4
    assert lib.j2kcuda_dwt_forward_noalloc(src.data_ptr() + 4, dst.data_ptr(), w, 0, 0, w - 4, h, 2, 0, None) == L.ERR_PARAM  # alignment
