"""One chunk of the C3 workload, device-resident, a few passes: the command ncu profiles
(scripts/gpu_ncu_strip.sh).  argv: pairs [compact 0|1] [del ext]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W

pairs_n = int(sys.argv[1]) if len(sys.argv) > 1 else 1600000
compact = len(sys.argv) > 2 and sys.argv[2] == "1"
d = float(sys.argv[3]) if len(sys.argv) > 3 else 10.0
e = float(sys.argv[4]) if len(sys.argv) > 4 else 2.0
q, t = W.dna_read_pairs_torch(0, pairs_n, 150)
pairs = ab.PackedPairs.uniform(q, t)
src = pairs.pack2bit() if compact else pairs
db = ab.DeviceBatch(N.LOCAL, src, ab.dna_matrix(5, -4), d, e, compact=compact)
ms, fill = db.run_timed(3, profile_fill=True)
print("pairs %d ms/step %.3f fill %.3f GCUPS %.1f" % (pairs_n, ms / 3, fill / 3, pairs_n * 22500.0 / (ms / 3e3) / 1e9))
db.close()
