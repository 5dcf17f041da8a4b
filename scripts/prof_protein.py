"""A protein batch on the INT32 warp kernel: the command ncu profiles.  argv: pairs [del ext]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W
P = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 11
e = int(sys.argv[3]) if len(sys.argv) > 3 else 2
pp = W.protein_pairs(0, P, 270, 330)
cells = float((np.diff(pp.q_off).astype(np.float64) * np.diff(pp.t_off)).sum())
for so in (False, True):
    db = ab.DeviceBatch(N.LOCAL, pp, ab.get_blosum62(), d, e, score_only=so)
    db.run_timed(1, profile_fill=False)
    ms, fill = db.run_timed(3, profile_fill=True)
    print("%d protein pairs d=%d e=%d %s: %.3f ms per pass (fill %.3f) = %.1f GCUPS" %
          (P, d, e, "score-only" if so else "with traceback", ms / 3, fill / 3, cells / (ms / 3e3) / 1e9))
    db.close()
