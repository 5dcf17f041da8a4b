"""Intra-pair wavefront kernel (BASELINE config 4 shapes) against the lean oracle."""
import numpy as np
import pytest

import oracle_lib as O

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _long_pair(n, seed):
    q, t = W.dna_long_pair(n, seed=seed)
    return q, t


@pytest.mark.parametrize("n,mode,d,e", [(4200, "global", 8, 8), (4200, "global", 10, 2), (4200, "local", 6, 6),
                                        (9000, "global", 8, 8), (9000, "local", 7, 7)])
def test_wavefront_vs_lean_oracle(n, mode, d, e):
    q, t = _long_pair(n, seed=4 + n)
    M = ab.dna_matrix(5, -4)
    ref = O.align_pair(mode, q, t, float(d), float(e), M, lean=True)
    assert ref["status"] == 0
    cls = ab.SimpleGlobalAligner if mode == "global" else ab.SimpleLocalAligner
    r = cls.from_seqs(q, t, symbols=ab.DNA).perform_alignment(d, e, M, want_matrices=False)
    assert r.numeric == N.NUM_I32
    assert r.alignment.f == ref["f"] and r.corner_score == ref["corner"]
    assert r.alignment.coords == ref["coords"]
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()


def test_wavefront_rectangular_and_in_ragged_batch():
    rng = np.random.default_rng(3)
    big_q = rng.integers(0, 4, 7001).astype(np.uint8)
    big_t = np.concatenate([big_q[:3000], rng.integers(0, 4, 333).astype(np.uint8), big_q[3000:5123]])
    qs = [rng.integers(0, 4, 80).astype(np.uint8), big_q, rng.integers(0, 4, 33).astype(np.uint8)]
    ts = [rng.integers(0, 4, 70).astype(np.uint8), big_t, rng.integers(0, 4, 90).astype(np.uint8)]
    pairs = ab.PackedPairs.from_lists(qs, ts)
    M = ab.dna_matrix(3, -2)
    for mode in ("global", "local"):
        res = ab.align_batch(N.GLOBAL if mode == "global" else N.LOCAL, pairs, M, 5, 5)
        for i in range(3):
            ref = O.align_pair(mode, qs[i], ts[i], 5.0, 5.0, M, lean=True)
            assert res.status[i] == 0 and res.f[i] == ref["f"] and res.corner[i] == ref["corner"]
            qa, ta = res.alignment(i)
            assert (qa == ref["q_aln"]).all() and (ta == ref["t_aln"]).all()
            assert tuple(int(v) for v in res.coords[i]) == ref["coords"][0] + ref["coords"][1]


def test_wavefront_protein_matrix():
    pairs = W.protein_pairs(0, 1, 4300, 4300)
    q, t = pairs.pair(0)
    M = ab.get_blosum62()
    ref = O.align_pair("global", q, t, 11.0, 1.0, M, lean=True)
    r = ab.SimpleGlobalAligner.from_seqs(q, t).perform_alignment(11, 1, M, want_matrices=False)
    assert r.corner_score == ref["corner"]
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()


@pytest.mark.parametrize("m,n", [(300, 5000), (5000, 300), (40000, 33), (33, 40000), (1500, 1501), (4097, 1023)])
@pytest.mark.parametrize("mode,d", [("global", 8), ("local", 6)])
def test_tile_pipeline_extreme_shapes(m, n, mode, d):
    """Shapes that stress the 4 x 4 tile pipeline and the window traceback: fewer tile columns than
    lanes, a single band, n not a multiple of 4, unrelated sequences (long gap runs leave the
    predicted window, global paths end on the border rows)."""
    rng = np.random.default_rng(m * 7 + n)
    q = rng.integers(0, 4, n).astype(np.uint8)
    t = rng.integers(0, 4, m).astype(np.uint8)
    if min(m, n) > 1000:  # a related stretch so that local mode has a long path
        k = min(m, n) // 2
        t[m - k:] = q[n - k:]
        t[m - k + 7] ^= 1
    M = ab.dna_matrix(5, -4)
    ref = O.align_pair(mode, q, t, float(d), float(d), M, lean=True)
    cls = ab.SimpleGlobalAligner if mode == "global" else ab.SimpleLocalAligner
    if ref["status"] != 0:
        with pytest.raises(Exception):
            cls.from_seqs(q, t, symbols=ab.DNA).perform_alignment(d, d, M, want_matrices=False)
        return
    r = cls.from_seqs(q, t, symbols=ab.DNA).perform_alignment(d, d, M, want_matrices=False)
    assert r.numeric == N.NUM_I32
    assert r.alignment.f == ref["f"] and r.corner_score == ref["corner"]
    assert r.alignment.coords == ref["coords"]
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()


def test_cooperative_walk_equals_lockstep_walk(monkeypatch):
    """The two long-pair tracebacks — the cooperative window walk (full direction matrix) and the
    lockstep walk (used group by group by the checkpointed traceback) — give the same alignment."""
    q, t = W.dna_long_pair(6000, seed=11)
    al = ab.SimpleGlobalAligner.from_seqs(q, t, symbols=ab.DNA)
    a = al.perform_alignment(8, 8, ab.dna_matrix(5, -4), want_matrices=False)
    monkeypatch.setenv("ALNR_SCRATCH_BUDGET", "100000000")  # one group holding every band
    b = al.perform_alignment(8, 8, ab.dna_matrix(5, -4), want_matrices=False)
    assert a.corner_score == b.corner_score
    assert (a.alignment.query == b.alignment.query).all() and (a.alignment.target == b.alignment.target).all()


@pytest.mark.parametrize("seed", range(24))
def test_tile_pipeline_and_cooperative_walk_fuzz(seed):
    """Random shapes, scoring schemes and degrees of relatedness (cheap gaps give long gap runs and
    paths far from the diagonal: wrong guesses, cut walks, windows left through the side, the
    one-lane fallback) against the lean oracle."""
    rng = np.random.default_rng(9000 + seed)
    m, n = (int(v) for v in rng.integers(1030, 3200, 2))
    q = rng.integers(0, 4, n).astype(np.uint8)
    t = rng.integers(0, 4, m).astype(np.uint8)
    kind = seed % 4
    if kind != 0:  # related stretches with substitutions and indels of growing length
        src, out, i = q, [], 0
        indel = [0.0, 0.01, 0.05, 0.15][kind]
        while i < len(src) and len(out) < m:
            r = rng.random()
            if r < indel / 2:
                i += int(rng.integers(1, 40))
            elif r < indel:
                out.extend(rng.integers(0, 4, int(rng.integers(1, 40))).tolist())
            else:
                out.append(int(src[i]) if rng.random() > 0.1 else int(rng.integers(0, 4)))
                i += 1
        t = np.array((out + rng.integers(0, 4, m).tolist())[:m], dtype=np.uint8)
    match, mis, d = [(5, -4, 8), (2, -3, 1), (1, -1, 2), (7, -5, 3)][(seed // 4) % 4]
    mode = "global" if seed % 3 else "local"
    M = ab.dna_matrix(match, mis)
    ref = O.align_pair(mode, q, t, float(d), float(d), M, lean=True)
    cls = ab.SimpleGlobalAligner if mode == "global" else ab.SimpleLocalAligner
    if ref["status"] != 0:
        with pytest.raises(Exception):
            cls.from_seqs(q, t, symbols=ab.DNA).perform_alignment(d, d, M, want_matrices=False)
        return
    r = cls.from_seqs(q, t, symbols=ab.DNA).perform_alignment(d, d, M, want_matrices=False)
    assert r.alignment.f == ref["f"] and r.corner_score == ref["corner"]
    assert r.alignment.coords == ref["coords"]
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()


def test_several_long_pairs_of_growing_size_in_one_batch():
    """Each long pair is its own chunk; the band-row scratch grows from pair to pair."""
    rng = np.random.default_rng(21)
    qs, ts = [], []
    for n in (1100, 1700, 2600, 1300):
        q = rng.integers(0, 4, n).astype(np.uint8)
        t = np.where(rng.random(n) < 0.85, q, rng.integers(0, 4, n)).astype(np.uint8)[: n - 37]
        qs.append(q)
        ts.append(t)
    pairs = ab.PackedPairs.from_lists(qs, ts)
    M = ab.dna_matrix(5, -4)
    for mode in ("global", "local"):
        res = ab.align_batch(N.GLOBAL if mode == "global" else N.LOCAL, pairs, M, 7, 7)
        for i in range(len(qs)):
            ref = O.align_pair(mode, qs[i], ts[i], 7.0, 7.0, M, lean=True)
            assert res.status[i] == 0 and res.f[i] == ref["f"] and res.corner[i] == ref["corner"]
            qa, ta = res.alignment(i)
            assert (qa == ref["q_aln"]).all() and (ta == ref["t_aln"]).all()


# ---- checkpointed traceback (aligner_b200/csrc/wave.cu::wave_ckpt_trace) ------------------------
# ALNR_SCRATCH_BUDGET (bytes) forces pairs whose direction matrix is larger onto the checkpointed
# path: score-only first pass that keeps the band rows, then re-fill + walk group by group.

@pytest.mark.parametrize("budget", [40_000, 300_000, 3_000_000])
@pytest.mark.parametrize("n,mode,d", [(5000, "global", 8), (5000, "local", 6), (9000, "global", 8)])
def test_checkpointed_traceback_vs_lean_oracle(monkeypatch, n, mode, d, budget):
    q, t = _long_pair(n, seed=40 + n)
    M = ab.dna_matrix(5, -4)
    ref = O.align_pair(mode, q, t, float(d), float(d), M, lean=True)
    monkeypatch.setenv("ALNR_SCRATCH_BUDGET", str(budget))  # 1 .. many bands per group
    cls = ab.SimpleGlobalAligner if mode == "global" else ab.SimpleLocalAligner
    r = cls.from_seqs(q, t, symbols=ab.DNA).perform_alignment(d, d, M, want_matrices=False)
    assert r.alignment.f == ref["f"] and r.corner_score == ref["corner"] and r.alignment.coords == ref["coords"]
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()


@pytest.mark.parametrize("m,n", [(300, 5000), (5000, 300), (40000, 33), (33, 40000), (4097, 1023)])
@pytest.mark.parametrize("mode,d", [("global", 8), ("local", 6)])
def test_checkpointed_traceback_extreme_shapes(monkeypatch, m, n, mode, d):
    """Unrelated sequences: global paths run along the border rows, local alignments are short and
    sit anywhere (most groups are skipped)."""
    rng = np.random.default_rng(m * 7 + n)
    q = rng.integers(0, 4, n).astype(np.uint8)
    t = rng.integers(0, 4, m).astype(np.uint8)
    M = ab.dna_matrix(5, -4)
    ref = O.align_pair(mode, q, t, float(d), float(d), M, lean=True)
    monkeypatch.setenv("ALNR_SCRATCH_BUDGET", "20000")
    cls = ab.SimpleGlobalAligner if mode == "global" else ab.SimpleLocalAligner
    r = cls.from_seqs(q, t, symbols=ab.DNA).perform_alignment(d, d, M, want_matrices=False)
    assert r.alignment.f == ref["f"] and r.corner_score == ref["corner"] and r.alignment.coords == ref["coords"]
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()


def test_checkpointed_traceback_equals_full_matrix_path_1e5(monkeypatch):
    """10^5 x 10^5 global: the checkpointed path (64 MB of direction scratch instead of 2.5 GB) returns
    the strings and the score of the full-matrix path."""
    q, t = W.dna_long_pair(100000, seed=4)
    M = ab.dna_matrix(5, -4)
    al = ab.SimpleGlobalAligner.from_seqs(q, t, symbols=ab.DNA)
    l0 = N.lib().alnr_kernel_launch_count()
    full = al.perform_alignment(8, 8, M, want_matrices=False)
    l1 = N.lib().alnr_kernel_launch_count()
    monkeypatch.setenv("ALNR_SCRATCH_BUDGET", str(64 << 20))
    ck = al.perform_alignment(8, 8, M, want_matrices=False)
    # 2.5 GB of directions in 64 MB groups: one re-fill + one walk launch per group
    assert N.lib().alnr_kernel_launch_count() - l1 > (l1 - l0) + 40
    assert ck.corner_score == full.corner_score and ck.alignment.coords == full.alignment.coords
    assert (ck.alignment.query == full.alignment.query).all() and (ck.alignment.target == full.alignment.target).all()


def test_checkpointed_pair_inside_a_batch(monkeypatch):
    rng = np.random.default_rng(5)
    q1, t1 = _long_pair(6000, seed=77)
    q2, t2 = _long_pair(5000, seed=78)
    qs = [rng.integers(0, 4, 60).astype(np.uint8), q1, q2]
    ts = [rng.integers(0, 4, 70).astype(np.uint8), t1, t2]
    pairs = ab.PackedPairs.from_lists(qs, ts)
    M = ab.dna_matrix(5, -4)
    ref = ab.align_batch(N.GLOBAL, pairs, M, 8, 8)
    monkeypatch.setenv("ALNR_SCRATCH_BUDGET", "500000")
    res = ab.align_batch(N.GLOBAL, pairs, M, 8, 8)
    cr = ab.align_batch(N.GLOBAL, pairs, M, 8, 8, compact=True).expand()
    for i in range(3):
        assert res.corner[i] == ref.corner[i]
        for got in (res, cr):
            qa, ta = got.alignment(i)
            qb, tb = ref.alignment(i)
            assert (qa == qb).all() and (ta == tb).all()


# ---- difference-recurrence kernel (kernels_diff.cuh): global mode, DNA, small integer scores ------

@pytest.mark.parametrize("m,n", [(128, 128), (129, 4001), (4000, 131), (2500, 2500), (3333, 5100), (6007, 2049)])
@pytest.mark.parametrize("d,e,mat", [(8, 8, (5, -4)), (10, 2, (5, -4)), (3, 1, (1, -1)), (24, 24, (10, -9))])
def test_diff_kernel_vs_lean_oracle_and_tile_kernel(monkeypatch, m, n, d, e, mat):
    """Every band / lane / tile boundary case of the skewed sweep (rows and columns not multiples of
    4, 8, 64 or 128, one band, many bands), penalties with del > ext (cell (1,1) pays del) and the
    +1/-1 scheme full of ties — against the lean oracle and against the tile pipeline."""
    rng = np.random.default_rng(m * 31 + n + d)
    q = rng.integers(0, 4, n).astype(np.uint8)
    t = rng.integers(0, 4, m).astype(np.uint8)
    k = min(m, n)
    keep = rng.random(k) < 0.85
    t[:k] = np.where(keep, q[:k], t[:k])  # related prefix, unrelated tail: diagonal runs and long gaps
    M = ab.dna_matrix(*mat)
    ref = O.align_pair("global", q, t, float(d), float(e), M, lean=True)
    al = ab.SimpleGlobalAligner.from_seqs(q, t, symbols=ab.DNA)
    monkeypatch.delenv("ALNR_NO_DIFF", raising=False)
    r = al.perform_alignment(d, e, M, want_matrices=False)
    assert r.corner_score == ref["corner"] and r.alignment.f == 0.0 and r.alignment.coords == ref["coords"]
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()
    monkeypatch.setenv("ALNR_NO_DIFF", "1")
    r2 = al.perform_alignment(d, e, M, want_matrices=False)
    assert r2.corner_score == r.corner_score
    assert (r2.alignment.query == r.alignment.query).all() and (r2.alignment.target == r.alignment.target).all()


# ---- stateful local pairs on the tile pipeline (del != ext; simple/mod.rs:201-205) --------------------

def _stateful_case(kind, n, seed):
    rng = np.random.default_rng(seed)
    if kind == "similar":
        return _long_pair(n, seed)
    if kind == "unrelated":
        return rng.integers(0, 4, n).astype(np.uint8), rng.integers(0, 4, n - 37).astype(np.uint8)
    # "island": unrelated flanks around a shared core — the best cell lies far from every border
    core = rng.integers(0, 4, n // 3).astype(np.uint8)
    q = np.concatenate([rng.integers(0, 4, n // 3).astype(np.uint8), core, rng.integers(0, 4, n // 4).astype(np.uint8)])
    t = np.concatenate([rng.integers(0, 4, n // 5).astype(np.uint8), core, rng.integers(0, 4, n // 3).astype(np.uint8)])
    return q, t


@pytest.mark.parametrize("kind,n,match,mis,d,e", [
    ("similar", 4200, 5, -4, 10, 2), ("similar", 9000, 5, -4, 11, 2), ("island", 6000, 5, -4, 10, 2),
    ("unrelated", 3000, 5, -4, 10, 2),
    # small scores: row m holds many zeros, the penalty map of row 1 needs several passes to settle
    ("unrelated", 3000, 2, -2, 4, 2), ("island", 5000, 1, -1, 3, 1), ("unrelated", 2600, 1, -1, 2, 1)])
def test_stateful_local_long_pair_vs_lean_oracle(kind, n, match, mis, d, e, monkeypatch):
    q, t = _stateful_case(kind, n, seed=11 + n + d)
    M = ab.dna_matrix(match, mis)
    ref = O.align_pair("local", q, t, float(d), float(e), M, lean=True)
    monkeypatch.setenv("ALNR_WAVE_STATS", "1")
    pairs = ab.PackedPairs.from_lists([q], [t])
    db = ab.DeviceBatch(N.LOCAL, pairs, M, d, e)
    db.run()
    passes = db.rerun_count
    res = db.download()
    db.close()
    assert 1 <= passes <= 8, passes            # it ran on the tile pipeline
    assert res.status[0] == ref["status"]
    if ref["status"] == 0:
        assert res.f[0] == ref["f"] and res.corner[0] == ref["corner"]
        qa, ta = res.alignment(0)
        assert (qa == ref["q_aln"]).all() and (ta == ref["t_aln"]).all()
        assert tuple(int(v) for v in res.coords[0]) == ref["coords"][0] + ref["coords"][1]
    # the warp kernel (the path such pairs took before) gives the same bytes
    monkeypatch.setenv("ALNR_NO_WAVE_STATEFUL", "1")
    old = ab.align_batch(N.LOCAL, pairs, M, d, e)
    assert old.status[0] == res.status[0] and old.f[0] == res.f[0] and old.corner[0] == res.corner[0]
    qo, to = old.alignment(0)
    qn, tn = res.alignment(0)
    assert (qo == qn).all() and (to == tn).all() and (old.coords == res.coords).all()


def test_stateful_local_score_only_and_in_batch():
    rng = np.random.default_rng(8)
    big_q, big_t = _stateful_case("island", 7000, 5)
    qs = [rng.integers(0, 4, 60).astype(np.uint8), big_q, rng.integers(0, 4, 45).astype(np.uint8)]
    ts = [rng.integers(0, 4, 70).astype(np.uint8), big_t, rng.integers(0, 4, 90).astype(np.uint8)]
    pairs = ab.PackedPairs.from_lists(qs, ts)
    M = ab.dna_matrix(3, -2)
    res = ab.align_batch(N.LOCAL, pairs, M, 6, 1)
    sc = ab.score_batch(N.LOCAL, pairs, M, 6, 1)
    for i in range(3):
        ref = O.align_pair("local", qs[i], ts[i], 6.0, 1.0, M, lean=True)
        assert res.status[i] == ref["status"] == sc.status[i]
        assert res.f[i] == ref["f"] == sc.f[i] and res.corner[i] == ref["corner"] == sc.corner[i]
        qa, ta = res.alignment(i)
        assert (qa == ref["q_aln"]).all() and (ta == ref["t_aln"]).all()


@pytest.mark.parametrize("m,n", [(40, 140000), (140000, 36), (131, 33000), (4097, 1025)])
def test_stateful_local_extreme_shapes(m, n, monkeypatch):
    """Few columns, few rows, sizes around band / tile / map-word boundaries."""
    rng = np.random.default_rng(m + n)
    q = rng.integers(0, 4, n).astype(np.uint8)
    t = rng.integers(0, 4, m).astype(np.uint8)
    k = min(m, n)
    t[:k] = np.where(rng.random(k) < 0.8, q[:k], t[:k])
    M = ab.dna_matrix(2, -1)
    ref = O.align_pair("local", q, t, 3.0, 1.0, M, lean=True)
    monkeypatch.setenv("ALNR_WAVE_STATS", "1")
    pairs = ab.PackedPairs.from_lists([q], [t])
    db = ab.DeviceBatch(N.LOCAL, pairs, M, 3, 1)
    db.run()
    passes = db.rerun_count
    res = db.download()
    db.close()
    assert 1 <= passes <= 8, passes
    assert res.status[0] == ref["status"] == 0
    assert res.f[0] == ref["f"] and res.corner[0] == ref["corner"]
    qa, ta = res.alignment(0)
    assert (qa == ref["q_aln"]).all() and (ta == ref["t_aln"]).all()


def test_stateful_local_protein_long_pair():
    pairs = W.protein_pairs(0, 1, 4300, 4300)
    q, t = pairs.pair(0)
    M = ab.get_blosum62()
    ref = O.align_pair("local", q, t, 11.0, 1.0, M, lean=True)
    r = ab.SimpleLocalAligner.from_seqs(q, t).perform_alignment(11, 1, M, want_matrices=False)
    assert r.alignment.f == ref["f"] and r.corner_score == ref["corner"] and r.alignment.coords == ref["coords"]
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()
