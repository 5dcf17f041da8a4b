"""INT32 MODE 1 of fill_warp_kernel (static penalty, compile-time): find the smallest failing case.
ALNR_WARP_INT_MODE1=1 routes INT32 local del == ext batches to the MODE 1 instantiation."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_lib as O
import aligner_b200 as ab
from aligner_b200 import _native as N


def ragged(rng, count, lo, hi, V):
    qs, ts = [], []
    for _ in range(count):
        n = int(rng.integers(lo, hi + 1))
        q = rng.integers(0, V, n).astype(np.uint8)
        t = q.copy()
        mut = rng.random(n) < 0.25
        t[mut] = rng.integers(0, V, int(mut.sum()))
        qs.append(q)
        ts.append(t)
    return qs, ts


def main():
    only = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    M = ab.get_blosum62()
    bad_total = 0
    for hi in ([only] if only else [40, 90, 100, 150, 180, 220, 250, 280, 310, 340, 380, 410, 440, 470, 512]):
        rng = np.random.default_rng(hi)
        qs, ts = ragged(rng, 96 if not only else 8, max(24, hi - 60), hi, 24)
        pairs = ab.PackedPairs.from_lists(qs, ts)
        ref = O.align_batch("local", pairs, 5.0, 5.0, M, threads=4)
        res = ab.align_batch(N.LOCAL, pairs, M, 5, 5, numeric=N.NUM_I32)
        bad = np.nonzero((res.f != ref["f"]) | (res.coords != ref["coords"]).any(axis=1))[0]
        sbad = 0
        for i in range(pairs.count):
            L = int(ref["aln_len"][i])
            qa, ta = res.alignment(i)
            if len(qa) != L or (qa != ref["q_aln"][i, :L]).any() or (ta != ref["t_aln"][i, :L]).any():
                sbad += 1
        m_max = max(len(t) for t in ts)
        print("hi=%d R=%d pairs=%d wrong_f_or_coords=%d wrong_strings=%d" % (hi, max(2, (m_max + 31) // 32), pairs.count, len(bad), sbad))
        for i in bad[:4]:
            print("   pair %d n=%d m=%d f=%g ref=%g coords=%s ref=%s" % (i, len(qs[i]), len(ts[i]), res.f[i], ref["f"][i],
                                                                      res.coords[i].tolist(), ref["coords"][i].tolist()))
        bad_total += len(bad) + sbad
    print("TOTAL_BAD", bad_total)


if __name__ == "__main__":
    main()
