"""GPU parity of the score-only entry points (alnr_score_batch, SURVEY.md 8f row N2: the shuffle
batch of statistics::calculate_p_value, aligner-core/src/statistics/mod.rs:240-307) against the
oracle, in every numeric domain and kernel family."""
import numpy as np
import pytest

import oracle_lib as O

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import statistics as S
from aligner_b200 import workloads as W
from test_gpu_parity import MODE, _rand_uniform_pairs
from test_warp import _ragged

pytestmark = pytest.mark.gpu


def _check_scores(mode, pairs, d, e, M, numeric=N.NUM_AUTO):
    ref = O.align_batch(mode, pairs, float(d), float(e), M, threads=8, want_alignments=False)
    res = ab.score_batch(MODE[mode], pairs, M, d, e, numeric=numeric)
    want_status = np.array([O.STATUS_TO_ALNR[int(s)] for s in ref["status"]], dtype=np.int32)
    assert (res.status == want_status).all()
    ok = want_status == 0
    assert (res.f[ok] == ref["f"][ok]).all()
    assert (res.corner[ok] == ref["corner"][ok]).all()
    if mode == "local":  # coords = (.., sx + 1, .., sy + 1): simple/mod.rs:254-257
        assert (res.end_yx[ok, 0] == ref["coords"][ok, 3] - 1).all()
        assert (res.end_yx[ok, 1] == ref["coords"][ok, 1] - 1).all()
    else:
        assert (res.end_yx[ok, 0] == ref["coords"][ok, 3]).all()
        assert (res.end_yx[ok, 1] == ref["coords"][ok, 1]).all()
    return res


@pytest.mark.parametrize("mode,d,e", [("local", 10, 2), ("local", 6, 6), ("global", 10, 2), ("global", 4, 4)])
def test_strip_kernel_scores(mode, d, e):
    """uniform DNA reads -> packed 16-bit strip kernel, SCORE instantiations (R = 15 and 16)."""
    rng = np.random.default_rng(5)
    for n, m in [(150, 150), (64, 96), (33, 17)]:
        pairs = _rand_uniform_pairs(rng, 700, n, m)
        res = _check_scores(mode, pairs, d, e, ab.dna_matrix(5, -4), numeric=N.NUM_I16X2)
        assert res.status.shape == (700,)


def test_strip_kernel_scores_speculation_leftovers():
    """+1/-1, del 2, ext 1 on unrelated reads: re-run passes and the exact kernel, all without D."""
    rng = np.random.default_rng(6)
    pairs = _rand_uniform_pairs(rng, 900, 100, 100, related=False)
    _check_scores("local", pairs, 2, 1, ab.dna_matrix(1, -1), numeric=N.NUM_I16X2)


def test_scores_equal_alignment_call_on_config3_sample():
    pairs = W.dna_read_batch(3, 20000, n=150)
    M = ab.dna_matrix(5, -4)
    a = ab.align_batch(N.LOCAL, pairs, M, 10, 2, want_alignments=False)
    s = ab.score_batch(N.LOCAL, pairs, M, 10, 2)
    assert (a.f == s.f).all() and (a.corner == s.corner).all() and (a.status == s.status).all()


@pytest.mark.parametrize("numeric", [N.NUM_I32, N.NUM_F64])
def test_warp_and_exact_kernel_scores(numeric):
    rng = np.random.default_rng(7)
    pairs = _ragged(rng, 300, 24, 330, 24)
    for mode, d, e in [("local", 11, 2), ("local", 5, 5), ("global", 11, 1)]:
        _check_scores(mode, pairs, d, e, ab.get_blosum62(), numeric=numeric)
    tiny = _ragged(rng, 5000, 4, 20, 4, related=False)  # one pair per thread
    _check_scores("local", tiny, 3, 1, ab.dna_matrix(2, -3), numeric=numeric)


def test_long_pair_scores_band_kernel():
    rng = np.random.default_rng(8)
    q = rng.integers(0, 4, 1500).astype(np.uint8)
    t = np.where(rng.random(1500) < 0.85, q, rng.integers(0, 4, 1500)).astype(np.uint8)[:1300]
    pairs = ab.PackedPairs.from_lists([q], [t])
    for mode, d, e in [("local", 6, 6), ("global", 10, 2)]:
        _check_scores(mode, pairs, d, e, ab.dna_matrix(5, -4))


def test_bad_pairs_keep_their_status():
    a = np.array([0, 1, 2, 3, 1], dtype=np.uint8)
    pairs = ab.PackedPairs.from_lists([a, a[:0], np.array([98], dtype=np.uint8), a], [a, a, a, a[:0]])
    _check_scores("local", pairs, 2, 1, ab.dna_matrix(1, -1))


def test_p_value_shuffle_batch():
    """The reference's own shape: one protein query against 4999 shuffled, shortened targets
    (statistics/mod.rs:255-292); scores and lengths as calculate_distribution_params gets them."""
    assert S.shuffle_count() == 4999
    rng = np.random.default_rng(11)
    pp = W.protein_pairs(1, 1, 280, 320)
    query, target = pp.pair(0)
    targets = S.shuffled_targets(target, rng=rng)
    assert len(targets) == 4999
    assert all(len(target) - 6 <= len(t) <= len(target) for t in targets)
    assert all((np.sort(t) == np.sort(target[:len(t)])).all() for t in targets[:50])
    M = ab.get_blosum62()
    scores, lengths, res = ab.shuffle_scores(query, target, 123.0, 11, 2, M, targets=targets)
    assert scores.shape == (5000,) and lengths.shape == (5000,)
    assert scores[0] == 123.0 and lengths[0] == len(target)
    sub = np.arange(0, 4999, 7)
    pairs = ab.PackedPairs.from_lists([query] * len(sub), [targets[i] for i in sub])
    ref = O.align_batch("local", pairs, 11.0, 2.0, M, threads=8, want_alignments=False)
    assert (ref["status"] == 0).all()
    assert (scores[1:][sub] == ref["f"]).all()
    assert (lengths[1:][sub] == [len(targets[i]) for i in sub]).all()


def test_device_batch_score_only_handle():
    pairs = W.dna_read_batch(9, 4096, n=150)
    M = ab.dna_matrix(5, -4)
    db = ab.DeviceBatch(N.LOCAL, pairs, M, 10, 2, score_only=True)
    ms, fill = db.run_timed(2)
    got = db.download_scores()
    want = ab.score_batch(N.LOCAL, pairs, M, 10, 2)
    assert ms > 0 and fill > 0
    for k in ("f", "corner", "end_yx", "status"):
        assert (getattr(got, k) == getattr(want, k)).all()
    with pytest.raises(N.AlignerError):
        db.download()  # alignments were never produced
    db.close()
