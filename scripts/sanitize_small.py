"""One small launch of every kernel family, for compute-sanitizer (memcheck / racecheck / initcheck /
synccheck): strip (speculation + re-runs), warp INT32 / FP64 (stateful, multi-band), tile pipeline +
cooperative window walk, exact kernel, compact formats, heuristic search, PWM."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W

which = sys.argv[1:] or ["strip", "warp", "tile", "compact", "heur"]
M = ab.dna_matrix(5, -4)
rng = np.random.default_rng(1)
if "strip" in which:
    p = W.dna_read_batch(0, 512, n=150)
    r = ab.align_batch(N.LOCAL, p, M, 10, 2, numeric=N.NUM_I16X2)
    r = ab.align_batch(N.GLOBAL, p, M, 8, 8, numeric=N.NUM_I16X2)
    ab.score_batch(N.LOCAL, p, M, 10, 2, numeric=N.NUM_I16X2)
    print("strip ok", float(r.corner.sum()))
if "warp" in which:
    qs = [rng.integers(0, 24, int(rng.integers(40, 700))).astype(np.uint8) for _ in range(24)]
    ts = [np.where(rng.random(len(q)) < 0.7, q, rng.integers(0, 24, len(q))).astype(np.uint8) for q in qs]
    p = ab.PackedPairs.from_lists(qs, ts)
    B = ab.get_blosum62()
    for num in (N.NUM_I32, N.NUM_F64):
        r = ab.align_batch(N.LOCAL, p, B, 11, 2, numeric=num)
        r = ab.align_batch(N.LOCAL, p, B, 5, 5, numeric=num)
        r = ab.align_batch(N.GLOBAL, p, B, 11, 1, numeric=num)
    al = ab.SimpleLocalAligner.from_seqs(qs[0], ts[0], symbols=ab.Protein)
    al.perform_alignment(11, 2, B)
    print("warp ok", float(r.corner.sum()))
if "tile" in which:
    q, t = W.dna_long_pair(3000, seed=9)
    p = ab.PackedPairs.from_lists([q], [t])
    r = ab.align_batch(N.GLOBAL, p, M, 8, 8)
    r2 = ab.align_batch(N.LOCAL, p, M, 6, 6)
    print("tile ok", float(r.corner[0]), float(r2.f[0]))
if "compact" in which:
    p = W.dna_read_batch(0, 512, n=150)
    c = ab.align_batch(N.LOCAL, p.pack2bit(), M, 10, 2, numeric=N.NUM_I16X2, compact=True)
    qs = [rng.integers(0, 4, int(rng.integers(20, 200))).astype(np.uint8) for _ in range(64)]
    p = ab.PackedPairs.from_lists(qs, qs)
    c2 = ab.align_batch(N.LOCAL, p.pack2bit(), M, 10, 2, compact=True)
    print("compact ok", c.move_words(), c2.move_words())
if "heur" in which:
    pairs = W.protein_pairs(0, 6, 60, 90)
    mats = W.candidate_matrices(4, 24, seed=5)
    fr = W.pair_frequencies(pairs, 24)
    out = ab.matrix_search(pairs, mats, 5.0, 5.0, -1.0, 0.0, fr)
    print("heur ok", float(np.asarray(out[0]).sum()) if isinstance(out, tuple) else "done")
