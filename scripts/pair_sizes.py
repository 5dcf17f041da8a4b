"""Latency of one global static-penalty call against the pair size (which kernel family serves it)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, aligner_b200 as ab
from aligner_b200 import workloads as W
for n in (100, 200, 300, 450, 600, 800, 1000, 1400):
    q, t = W.dna_long_pair(n, seed=3)
    al = ab.SimpleGlobalAligner.from_seqs(q, t, symbols=ab.DNA)
    al.perform_alignment(8, 8, ab.dna_matrix(5, -4), want_matrices=False)
    ts = []
    for _ in range(30):
        t0 = time.perf_counter(); r = al.perform_alignment(8, 8, ab.dna_matrix(5, -4), want_matrices=False); ts.append(time.perf_counter() - t0)
    ts.sort()
    print("n=%5d cells=%8d median %.3f ms score %g" % (n, len(q) * len(t), ts[15] * 1e3, r.corner_score), flush=True)
