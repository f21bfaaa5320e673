# memory checker over small round trips (odd sizes, tiles, 9/7, cinema) -- compute-sanitizer memcheck
mkdir -p gpurun_out
cat > /tmp/san.py <<'PY'
import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np, j2k
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame
for (h, w, nc, prec, kw) in [(97, 131, 3, 8, dict(mct=1)), (130, 257, 4, 16, dict(mct=1)), (40, 100, 3, 16, dict(mct=1, tile_w=48, tile_h=48)),
                             (1, 1, 1, 8, dict(numresolution=1)), (64, 64, 1, 8, dict(numresolution=1)), (200, 333, 3, 12, dict(irreversible=1, mct=1, rates=[8.0])),
                             (135, 256, 3, 12, dict(irreversible=1, mct=1, profile=L.PROFILE_CINEMA_4K)), (33, 100, 1, 1, dict(numresolution=3)),
                             (300, 500, 3, 8, dict(mct=1, cblk_w=32, cblk_h=16, prog_order=2))]:
    for kind in ("noise", "smooth"):
        img = make_frame(h, w, nc, prec, kind, 3)
        cs = j2k.encode(img, prec, **kw)
        out = j2k.decode(cs)
        if not kw.get("irreversible"):
            assert np.array_equal(out, img)
        for r in (1, 2):
            if r < kw.get("numresolution", 6):
                j2k.decode(cs, reduce=r)
print("sanitizer workload done")
PY
compute-sanitizer --tool memcheck --error-exitcode 9 python /tmp/san.py > gpurun_out/p_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -6 gpurun_out/p_memcheck.log
