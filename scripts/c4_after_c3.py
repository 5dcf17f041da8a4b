"""C4 10^5 x 10^5 device-resident timing before and after a C3 strip-kernel batch ran in the same
process: do the strip kernel's evict_last boundary rows linger in L2 and slow the tile pipeline?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W

M = ab.dna_matrix(5, -4)
q, t = W.dna_long_pair(100000, seed=4)
pp = ab.PackedPairs.from_lists([q], [t])


def c4(tag):
    db = ab.DeviceBatch(N.GLOBAL, pp, M, 8, 8)
    db.run_timed(2, profile_fill=False)
    ms, _ = db.run_timed(3, profile_fill=False)
    _, fill = db.run_timed(3, profile_fill=True)
    db.close()
    print("%s: C4 %.2f ms per call, fill %.2f ms" % (tag, ms / 3, fill / 3), flush=True)


c4("fresh process")
qq, tt = W.dna_read_pairs_torch(0, 3000000, 150)
reads = ab.PackedPairs.uniform(qq, tt)
db = ab.DeviceBatch(N.LOCAL, reads, M, 10, 2)
db.run_timed(3, profile_fill=False)
db.close()
c4("after 3 C3 passes")
c4("again")
