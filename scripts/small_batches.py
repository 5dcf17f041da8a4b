"""Which integer kernel family serves a batch of 150 bp reads best, against the batch size:
the packed 16-bit strip kernel (two pairs per thread) or the INT32 warp kernel (one warp per pair)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W
M = ab.dna_matrix(5, -4)
for cnt in (1, 64, 1024, 8192, 32768, 131072, 524288):
    pairs = W.dna_read_batch(0, cnt, n=150)
    row = []
    for num in (N.NUM_I16X2, N.NUM_I32):
        ab.align_batch(N.LOCAL, pairs, M, 10, 2, numeric=num)
        ts = []
        for _ in range(5):
            t0 = time.perf_counter(); ab.align_batch(N.LOCAL, pairs, M, 10, 2, numeric=num); ts.append(time.perf_counter() - t0)
        row.append(min(ts) * 1e3)
    print("pairs %7d: strip i16x2 %.3f ms, warp i32 %.3f ms" % (cnt, row[0], row[1]), flush=True)
