"""Long LOCAL pairs with del != ext: the tile pipeline's speculate-and-repeat passes against the one-warp
path they took before (ALNR_NO_WAVE_STATEFUL=1).  argv: sizes."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W

M = ab.dna_matrix(5, -4)
os.environ["ALNR_WAVE_STATS"] = "1"
for n in [int(v) for v in sys.argv[1:]] or [20000, 100000]:
    q, t = W.dna_long_pair(n, seed=4)
    pairs = ab.PackedPairs.from_lists([q], [t])
    out = {}
    for tag, env in (("tile pipeline", None), ("one warp", "1")):
        if env: os.environ["ALNR_NO_WAVE_STATEFUL"] = env
        else: os.environ.pop("ALNR_NO_WAVE_STATEFUL", None)
        if tag == "one warp" and n > 50000:
            continue
        db = ab.DeviceBatch(N.LOCAL, pairs, M, 11, 2)
        db.run(); db.sync()
        if not env:
            ms, fill = db.run_timed(2, profile_fill=True)
            print("   run_timed: %.2f ms per call, fill %.2f ms" % (ms / 2, fill / 2))
            db2 = ab.DeviceBatch(N.LOCAL, pairs, M, 11, 11)
            db2.run_timed(1, profile_fill=False)
            ms, fill = db2.run_timed(2, profile_fill=True)
            print("   static (d = e = 11) run_timed: %.2f ms per call, fill %.2f ms" % (ms / 2, fill / 2))
            db2.close()
        t0 = time.perf_counter()
        db.run(); db.sync()
        dt = time.perf_counter() - t0
        passes = db.rerun_count
        res = db.download()
        db.close()
        out[tag] = (res.f[0], res.corner[0], tuple(res.coords[0]))
        print("%6d x %6d local d=11 e=2, %-13s: %9.2f ms per call (%6.1f GCUPS), passes %d, f %.0f" %
              (len(t), len(q), tag, dt * 1e3, len(q) * len(t) / dt / 1e9, passes, res.f[0]), flush=True)
    if len(out) == 2:
        assert out["tile pipeline"] == out["one warp"], out
