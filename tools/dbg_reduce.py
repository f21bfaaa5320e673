import sys, numpy as np
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import j2k, refgrok
from j2k_b200.frames import make_frame
img = make_frame(300, 421, 3, 8, "smooth", seed=6)
cs = j2k.encode(img, 8, mct=1, numresolution=5)
for reduce in (0,1,2):
    got = j2k.decode(cs, reduce=reduce)
    ref = refgrok.decode(cs, reduce=reduce)
    for c in range(3):
        d = np.argwhere(got[...,c].astype(np.int32) != ref[c])
        print('reduce',reduce,'comp',c,'shape',got.shape,'diffs',len(d), 'rows',(d[:,0].min(),d[:,0].max()) if len(d) else None,'cols',(d[:,1].min(),d[:,1].max()) if len(d) else None)
        if len(d): 
            y,x=d[0]; print('  first',y,x,got[y,x,c],ref[c][y,x], 'maxabs', np.abs(got[...,c].astype(np.int32)-ref[c]).max())
