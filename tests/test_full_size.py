"""Full-size runs of the BASELINE configs on the GPU, checked through size-independent properties
(the oracle only finishes samples of these sizes in seconds):
  C3  10^7 read pairs: slot/offset invariants, de-gapped strings == input ranges, the alignment
      re-scored from its strings == f (del == ext), score-only kernels == alignment kernels, the
      packed 16-bit strip kernel == the INT32 warp kernel, idempotence, oracle on a random sample;
  C4  10^5 x 10^5 global: de-gapped strings == the inputs, re-scored path == corner score;
  C5  2000 pairs x 100 matrices: fitness == the fixed-order blocked sums of the f grid.
ALNR_FULL_PAIRS shrinks C3 for quick local runs."""
import os

import numpy as np
import pytest

import oracle_lib as O

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W

pytestmark = pytest.mark.gpu

FULL_PAIRS = int(os.environ.get("ALNR_FULL_PAIRS", "10000000"))
BLANK = 98


@pytest.fixture(scope="module")
def reads():
    q, t = W.dna_read_pairs_torch(0, FULL_PAIRS, n=150)
    # the GPU generator is the bench's; it must produce the documented pairs
    q0, t0 = W.dna_read_pairs(0, 2000, n=150)
    assert (q[:2000] == q0).all() and (t[:2000] == t0).all()
    return ab.PackedPairs.uniform(q, t)


def _subset(pairs, idx):
    n, m = int(pairs.q_off[1]), int(pairs.t_off[1])
    q = pairs.q.reshape(-1, n)[idx]
    t = pairs.t.reshape(-1, m)[idx]
    return ab.PackedPairs.uniform(q, t)


def _slot_invariants(pairs, res):
    cap = pairs.aln_capacity()
    ln = res.aln_len.astype(np.int64)
    off = res.aln_off.astype(np.int64)
    n, m = int(pairs.q_off[1]), int(pairs.t_off[1])
    assert (ln >= 1).all() and (ln <= n + m + 1).all()
    order = np.argsort(off, kind="stable")
    so, sl = off[order], ln[order]
    assert so[0] >= 0 and so[-1] + sl[-1] <= cap
    assert (so[:-1] + sl[:-1] <= so[1:]).all()  # alignments never overlap


def _check_strings(pairs, res, idx, M, d, rescore):
    """de-gapped strings == the input ranges the coords name (+ the duplicated seed column,
    simple/mod.rs:213-216); with del == ext the strings re-score to f."""
    n = int(pairs.q_off[1])
    q2 = pairs.q.reshape(-1, n)
    t2 = pairs.t.reshape(-1, int(pairs.t_off[1]))
    for i in idx:
        qa, ta = res.alignment(i)
        x0, x1, y0, y1 = (int(v) for v in res.coords[i])
        assert ((qa != BLANK) | (ta != BLANK)).all()
        assert (qa[qa != BLANK] == np.append(q2[i, x0 - 1:x1 - 1], q2[i, x1 - 2])).all(), i
        assert (ta[ta != BLANK] == np.append(t2[i, y0 - 1:y1 - 1], t2[i, y1 - 2])).all(), i
        if rescore:
            body_q, body_t = qa[:-1], ta[:-1]  # the last column is the duplicated seed
            both = (body_q != BLANK) & (body_t != BLANK)
            s = M[body_t[both].astype(np.intp), body_q[both].astype(np.intp)].sum() - d * float((~both).sum())
            assert s == res.f[i], (i, s, res.f[i])


def test_config3_full_size_stateful_penalty(reads):
    M = ab.dna_matrix(5, -4)
    P = reads.count
    res = ab.align_batch(N.LOCAL, reads, M, 10, 2)
    assert (res.status == 0).all()
    _slot_invariants(reads, res)
    rng = np.random.default_rng(1)
    # SURVEY.md 8d / BASELINE.md: parity on a 10^5-pair subsample against the oracle, all outputs
    sample = np.sort(rng.choice(P, size=min(P, 100000), replace=False))
    _check_strings(reads, res, sample[:3000], M, 10, rescore=False)
    # oracle on the sample
    sub = _subset(reads, sample)
    ref = O.align_batch("local", sub, 10.0, 2.0, M, threads=os.cpu_count() or 8)
    assert (ref["status"] == 0).all()
    assert (res.f[sample] == ref["f"]).all() and (res.corner[sample] == ref["corner"]).all()
    assert (res.coords[sample] == ref["coords"]).all() and (res.aln_len[sample] == ref["aln_len"]).all()
    for k, i in enumerate(sample):  # every aligned string of the sample
        qa, ta = res.alignment(i)
        assert (qa == ref["q_aln"][k, :len(qa)]).all() and (ta == ref["t_aln"][k, :len(ta)]).all(), i
    # the score-only kernels agree with the alignment kernels on every pair
    sc = ab.score_batch(N.LOCAL, reads, M, 10, 2)
    assert (sc.f == res.f).all() and (sc.corner == res.corner).all() and (sc.status == 0).all()
    assert (sc.end_yx[:, 0].astype(np.uint64) == res.coords[:, 3] - 1).all()
    assert (sc.end_yx[:, 1].astype(np.uint64) == res.coords[:, 1] - 1).all()
    # idempotence: a second call gives the same bytes
    again = ab.align_batch(N.LOCAL, reads, M, 10, 2)
    for k in ("f", "corner", "coords", "aln_len", "aln_off"):
        assert (getattr(again, k) == getattr(res, k)).all(), k
    tot = int(res.aln_off.max() + 200)
    assert (again.q_aln[:tot] == res.q_aln[:tot]).all() and (again.t_aln[:tot] == res.t_aln[:tot]).all()
    # two kernel families: packed 16-bit strip kernel vs INT32 warp kernel on a contiguous slice
    lo = P // 3
    sl = np.arange(lo, min(P, lo + 1000000))
    part = _subset(reads, sl)
    i32 = ab.align_batch(N.LOCAL, part, M, 10, 2, numeric=N.NUM_I32)
    assert (i32.f == res.f[sl]).all() and (i32.coords == res.coords[sl]).all()
    assert (i32.aln_len == res.aln_len[sl]).all()
    for k in range(0, len(sl), 997):
        a, b = i32.alignment(k), res.alignment(int(sl[k]))
        assert (a[0] == b[0]).all() and (a[1] == b[1]).all()


def test_config3_full_size_static_penalty_rescores(reads):
    M = ab.dna_matrix(5, -4)
    res = ab.align_batch(N.LOCAL, reads, M, 6, 6)
    assert (res.status == 0).all()
    _slot_invariants(reads, res)
    rng = np.random.default_rng(2)
    sample = np.sort(rng.choice(reads.count, size=min(reads.count, 5000), replace=False))
    _check_strings(reads, res, sample, M, 6.0, rescore=True)
    glob = ab.align_batch(N.GLOBAL, reads, M, 6, 6, want_alignments=False)
    assert (glob.f == 0.0).all() and (glob.coords == np.array([1, 150, 1, 150], dtype=np.uint64)).all()
    sub = _subset(reads, sample)
    ref = O.align_batch("global", sub, 6.0, 6.0, M, threads=os.cpu_count() or 8, want_alignments=False)
    assert (glob.corner[sample] == ref["corner"]).all() and (glob.aln_len[sample] == ref["aln_len"]).all()


def test_config4_full_size_global_path():
    n = 100000
    q, t = W.dna_long_pair(n, seed=4)
    M = ab.dna_matrix(5, -4)
    al = ab.SimpleGlobalAligner.from_seqs(q, t, symbols=ab.DNA)
    r = al.perform_alignment(8, 8, M, want_matrices=False)
    qa, ta = r.alignment.query, r.alignment.target
    assert r.alignment.coords == ((1, len(q)), (1, len(t))) and r.alignment.f == 0.0
    assert ((qa != BLANK) | (ta != BLANK)).all()
    # the walk covers both sequences completely; the last column is the duplicated seed (:102-105)
    assert (qa[qa != BLANK] == np.append(q, q[-1])).all()
    assert (ta[ta != BLANK] == np.append(t, t[-1])).all()
    bq, bt = qa[:-1], ta[:-1]
    both = (bq != BLANK) & (bt != BLANK)
    s = M[bt[both].astype(np.intp), bq[both].astype(np.intp)].sum() - 8.0 * float((~both).sum())
    assert s == r.corner_score
    again = al.perform_alignment(8, 8, M, want_matrices=False)
    assert (again.alignment.query == qa).all() and (again.alignment.target == ta).all()
    sc = ab.score_batch(N.GLOBAL, ab.PackedPairs.from_lists([q], [t]), M, 8, 8)
    assert sc.corner[0] == r.corner_score
    loc = ab.SimpleLocalAligner.from_seqs(q, t, symbols=ab.DNA).perform_alignment(8, 8, M, want_matrices=False)
    lq, lt = loc.alignment.query[:-1], loc.alignment.target[:-1]
    both = (lq != BLANK) & (lt != BLANK)
    assert M[lt[both].astype(np.intp), lq[both].astype(np.intp)].sum() - 8.0 * float((~both).sum()) == loc.alignment.f


def test_tile_pipeline_more_bands_than_resident_warps():
    """600 000 target rows = 4688 bands, more than the grid holds at once (148 SMs x 28 blocks):
    blocks take a second band after their first.  Checked through the size-independent properties
    (the path covers both sequences, re-scores to the corner score) and against the score-only call."""
    rng = np.random.default_rng(77)
    m, n = 600000, 1800
    t = rng.integers(0, 4, m).astype(np.uint8)
    q = t[rng.integers(0, m - 2 * n)::333][:n].copy()
    M = ab.dna_matrix(5, -4)
    r = ab.SimpleGlobalAligner.from_seqs(q, t, symbols=ab.DNA).perform_alignment(8, 8, M, want_matrices=False)
    qa, ta = r.alignment.query, r.alignment.target
    assert (qa[qa != BLANK] == np.append(q, q[-1])).all()
    assert (ta[ta != BLANK] == np.append(t, t[-1])).all()
    bq, bt = qa[:-1], ta[:-1]
    both = (bq != BLANK) & (bt != BLANK)
    # cell (1, 1) pays del == ext here, and the corner rows use -(k+1)*del (simple/mod.rs:59-70): with
    # del == ext a path that never runs along row 0 / column 0 to the corner re-scores exactly
    s = M[bt[both].astype(np.intp), bq[both].astype(np.intp)].sum() - 8.0 * float((~both).sum())
    sc = ab.score_batch(N.GLOBAL, ab.PackedPairs.from_lists([q], [t]), M, 8, 8)
    assert sc.corner[0] == r.corner_score
    assert s == r.corner_score or s - 8.0 == r.corner_score  # -(m+1)*del corner quirk when the path ends on column 0


def test_config5_full_size_fitness_is_the_blocked_sum():
    P, K = 2000, 100
    pairs = W.protein_pairs(0, P, 270, 330)
    mats = W.candidate_matrices(K)
    freqs = W.pair_frequencies(pairs, 24)
    fitness, best_k, best_f, grid, sgrid = ab.matrix_search(pairs, mats, 5, 5, -1.0, 576.0, freqs)
    assert grid.shape == (P, K) and (sgrid == 0).sum() > 0.9 * P * K
    masked = np.where(sgrid == 0, grid, 0.0)
    want = np.zeros(K)
    for b0 in range(0, P, N.FITNESS_BLOCK):  # alnr_fitness_block_sums order
        blk = np.zeros(K)
        for i in range(b0, min(P, b0 + N.FITNESS_BLOCK)):
            blk = blk + masked[i]
        want = want + blk
    assert (fitness == want).all()
    ok = np.where(sgrid == 0, grid, -np.inf)
    assert (best_k == ok.argmax(axis=1)).all() and (best_f == ok.max(axis=1)).all()
    # a sample of items against the oracle's heuristic loop
    rng = np.random.default_rng(3)
    pi = rng.integers(0, P, 48).astype(np.uint32)
    mi = rng.integers(0, K, 48).astype(np.uint32)
    for p, k in zip(pi, mi):
        qq, tt = pairs.pair(int(p))
        ref = O.heuristic(qq, tt, 5.0, 5.0, mats[k], -1.0, 576.0, freqs[p])
        if ref["status"] == 0:
            assert sgrid[p, k] == 0 and grid[p, k] == ref["f"], (p, k)
        else:
            assert sgrid[p, k] != 0
