"""One encode + decode of a frame (for ncu captures): python tools/roundtrip_once.py H W NC PREC [kind] [reps]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import j2k
from j2k_b200.frames import make_frame
h, w, nc, prec = [int(v) for v in sys.argv[1:5]]
kind = sys.argv[5] if len(sys.argv) > 5 else "smooth"
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 2
img = make_frame(h, w, nc, prec, kind, 0)
for _ in range(reps):
    cs = j2k.encode(img, prec, mct=1 if nc >= 3 else 0)
    out = j2k.decode(cs)
assert np.array_equal(out, img)
print("ok", len(cs))
