"""One long global pair, a few passes: the command ncu profiles for the long-pair kernels.
argv: m n"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import aligner_b200 as ab
from aligner_b200 import _native as N
m = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
rng = np.random.default_rng(1)
q = rng.integers(0, 4, n).astype(np.uint8)
t = rng.integers(0, 4, m).astype(np.uint8)
k = min(m, n)
t[:k] = np.where(rng.random(k) < 0.85, q[:k], t[:k])
db = ab.DeviceBatch(N.GLOBAL, ab.PackedPairs.from_lists([q], [t]), ab.dna_matrix(5, -4), 8, 8)
ms, fill = db.run_timed(2, profile_fill=True)
print("%d x %d: %.3f ms per call, fill %.3f ms" % (m, n, ms / 2, fill / 2))
db.close()
