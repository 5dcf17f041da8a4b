"""Fill time of the long-pair pipeline kernels (kernels_tile.cuh; ALNR_BAND_COLUMN=1 selects the
first-generation column kernel of kernels_band.cuh): m rows (m/128 bands) x n columns, global mode."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, aligner_b200 as ab
from aligner_b200 import _native as N
rng = np.random.default_rng(1)
column = os.environ.get("ALNR_BAND_COLUMN") is not None
shapes = [(128, 400000), (1024, 400000), (128 * 148, 100000), (128 * 592, 50000), (100000, 100000)]
for m, n in shapes:
    q = rng.integers(0, 4, n).astype(np.uint8)
    t = rng.integers(0, 4, m).astype(np.uint8)
    db = ab.DeviceBatch(N.GLOBAL, ab.PackedPairs.from_lists([q], [t]), ab.dna_matrix(), 8, 8)
    db.run_timed(1)
    ms, fill = db.run_timed(1)
    bands = (m + 127) // 128
    steps = (n + 31 + (bands - 1) * 64) if column else ((n + 3) // 4 + 31 + (bands - 1) * 39)
    print("m=%6d n=%6d bands=%4d whole %.3f ms, fill %.3f ms -> %.1f ns per step (%.0f cycles), %.1f GCUPS fill" % (
        m, n, bands, ms, fill, fill * 1e6 / steps, fill * 1e6 / steps * 1.965, m * n / fill / 1e6))
    db.close()
