"""One long pair through the tile pipeline (for ncu): m x n, global mode."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, aligner_b200 as ab
from aligner_b200 import _native as N
m, n = int(sys.argv[1]), int(sys.argv[2])
rng = np.random.default_rng(1)
q = rng.integers(0, 4, n).astype(np.uint8)
t = rng.integers(0, 4, m).astype(np.uint8)
db = ab.DeviceBatch(N.GLOBAL, ab.PackedPairs.from_lists([q], [t]), ab.dna_matrix(), 8, 8)
print(db.run_timed(1))
db.close()
