"""Localise a decode mismatch: per-tile / per-component mismatch counts for a small tiled frame (debug aid)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import j2k
from j2k_b200.frames import make_frame

def report(h, w, nc, prec, numres, tile, mct, kind="noise"):
    img = make_frame(h, w, nc, prec, kind, seed=11)
    cs = j2k.encode(img, prec, mct=mct, numresolution=numres, tile_w=tile, tile_h=tile)
    got = j2k.decode(cs)
    bad = got != img
    line = "h%d w%d nc%d prec%d numres%d tile%d mct%d: %d bad" % (h, w, nc, prec, numres, tile, mct, int(bad.sum()))
    if bad.any():
        for c in range(nc):
            ys, xs = np.nonzero(bad[..., c])
            if len(ys):
                line += " | c%d n=%d x[%d..%d] y[%d..%d]" % (c, len(ys), xs.min(), xs.max(), ys.min(), ys.max())
    print(line, flush=True)

for numres in (6, 5, 4, 3, 2, 1):
    report(40, 100, 3, 16, numres, 48, 1)
report(40, 100, 3, 16, 6, 48, 0)
report(40, 100, 1, 16, 6, 48, 0)
report(40, 100, 1, 8, 6, 48, 0)
report(40, 52, 1, 16, 6, 48, 0)
report(40, 50, 1, 16, 6, 48, 0)
report(40, 49, 1, 16, 6, 48, 0)
report(40, 4, 1, 16, 3, 0, 0)
report(40, 2, 1, 16, 2, 0, 0)
report(40, 100, 1, 16, 6, 32, 0)
report(40, 100, 1, 16, 6, 64, 0)
report(70, 40, 1, 16, 6, 48, 0)
