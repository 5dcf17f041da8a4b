import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, aligner_b200 as ab
from aligner_b200 import _native as N, workloads as W
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
q, t = W.dna_long_pair(n, seed=4)
pairs = ab.PackedPairs.from_lists([q], [t])
db = ab.DeviceBatch(N.GLOBAL, pairs, ab.dna_matrix(), 8, 8)
db.run_timed(1)
for _ in range(2):
    ms, fill = db.run_timed(1)
    print("total ms %.2f fill ms %.2f -> fill GCUPS %.1f" % (ms, fill, len(q)*len(t)/fill/1e6))
