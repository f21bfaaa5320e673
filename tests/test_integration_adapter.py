"""The reference-side binding (integration/j2k_cuda_codec.*: a j2k::Codec subclass over the C ABI) compiles
against the reference's own unchanged headers (src/common/j2k_codec.h, j2k_io.h, j2k_exception.h)."""
import os
import subprocess
import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
REF = "/root/reference/src/common"


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "j2k_codec.h")), reason="reference headers not present")
def test_adapter_compiles_against_reference_headers():
    cmd = ["g++", "-std=c++11", "-fsyntax-only", "-Wall", "-include", "cstddef", "-include", "cstdint", "-include", "cstring",
           "-I" + REF, "-I/root/reference/ext/Little-CMS/include", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "integration", "j2k_cuda_codec.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
