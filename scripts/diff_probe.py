"""Fill time of the long-pair kernels for shapes that separate the step time (one band) from the
band-to-band hand-off (many bands).  ALNR_NO_DIFF=1 selects the tile pipeline."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import aligner_b200 as ab
from aligner_b200 import _native as N

M = ab.dna_matrix(5, -4)
rng = np.random.default_rng(1)
import json
shapes = json.loads(os.environ.get("PROBE_SHAPES", "[[128,100000],[1280,100000],[12800,100000],[100000,12800],[100000,100000]]"))
for m, n in shapes:
    q = rng.integers(0, 4, n).astype(np.uint8)
    t = rng.integers(0, 4, m).astype(np.uint8)
    k = min(m, n)
    t[:k] = np.where(rng.random(k) < 0.85, q[:k], t[:k])
    db = ab.DeviceBatch(N.GLOBAL, ab.PackedPairs.from_lists([q], [t]), M, 8, 8)
    db.run_timed(2, profile_fill=False)
    _, fill = db.run_timed(3, profile_fill=True)
    db.close()
    steps = n + m if not os.environ.get("ALNR_NO_DIFF") else (n + m) / 4
    print("%6d x %6d: fill %8.3f ms  -> %6.1f ns per (m+n) step, %7.1f GCUPS" % (m, n, fill / 3, fill / 3 * 1e6 / (n + m), m * n / (fill / 3e3) / 1e9), flush=True)
