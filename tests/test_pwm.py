"""GPU parity of the PWM aligner (aligner-core/src/pwm/mod.rs:29-126,
alignment.rs:45-79, heuristic/mod.rs:81-141) against the oracle restatement (oracle_pwm.cpp).
The reference holds no PWM test vectors: the oracle for this row is cross-checked only against the
independent Python restatement (tests/test_oracle.py)."""
import numpy as np
import pytest

import oracle_lib as O

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import pwm as P
from aligner_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _check_single(query, pwm, d, e, numeric=N.NUM_AUTO):
    ref = O.pwm_align(query, pwm, float(d), float(e))
    res = P.PWMAligner.from_seqs(query).perform_alignment(d, e, pwm, numeric=numeric)
    a = res.alignment
    assert a.f == ref["f"]
    assert a.coords == ref["coords"]
    assert (a.numbered == ref["numbered"]).all() and (a.query == ref["q_aln"]).all()
    assert a.dim == pwm.shape[1]
    assert (res.alignment_matrix == ref["H"]).all()
    assert (res.direction_matrix == ref["D"]).all()
    assert (a.get_frequency_matrix() == O.pwm_frequency_matrix(ref["numbered"], ref["q_aln"], pwm.shape[1])).all()
    return res


def _check_windows(seq, starts, lens, pwm, d, e, numeric=N.NUM_AUTO):
    ref = O.pwm_windows(seq, starts, lens, pwm, float(d), float(e))
    rs = P.ResidentSequence(seq)
    res = rs.align_windows(pwm, d, e, starts=starts, lens=lens, want_alignments=True, want_freq=True,
                           numeric=numeric)
    sc = rs.align_windows(pwm, d, e, starts=starts, lens=lens, numeric=numeric)
    rs.close()
    want = np.array([O.STATUS_TO_ALNR[int(s)] for s in ref["status"]], dtype=np.int32)
    assert (res.status == want).all() and (sc.status == want).all()
    ok = want == 0
    assert (res.f[ok] == ref["f"][ok]).all() and (sc.f[ok] == ref["f"][ok]).all()
    assert (res.coords[ok] == ref["coords"][ok]).all()
    assert (sc.coords[ok][:, [1, 3]] == ref["coords"][ok][:, [1, 3]]).all()  # score-only: arg-max only
    assert (res.aln_len[ok] == ref["aln_len"][ok]).all()
    L = pwm.shape[1]
    for i in np.nonzero(ok)[0]:
        a = res.alignment(i)
        n = int(ref["aln_len"][i])
        assert (a.numbered == ref["numbered"][i, :n]).all() and (a.query == ref["q_aln"][i, :n]).all(), i
        assert (res.freq[i] == O.pwm_frequency_matrix(a.numbered, a.query, L)).all(), i
    return res


@pytest.mark.parametrize("d,e", [(3, 1), (2, 2), (1.5, 0.25)])
def test_single_query_matrices_and_alignment(d, e):
    rng = np.random.default_rng(1)
    for m, L in [(40, 30), (330, 300), (7, 90), (90, 7), (1, 1), (600, 50)]:
        q = rng.integers(0, 4, m).astype(np.uint8)
        pwm = rng.integers(-2, 3, (4, L)).astype(np.float64)
        if d != int(d):
            pwm += rng.normal(0, 0.3, (4, L))
        _check_single(q, pwm, d, e)


def test_integer_and_fp64_domains_agree():
    rng = np.random.default_rng(2)
    q = rng.integers(0, 4, 200).astype(np.uint8)
    pwm = P.get_random_pwm(150, rng)
    a = _check_single(q, pwm, 2, 1, numeric=N.NUM_I32)
    b = _check_single(q, pwm, 2, 1, numeric=N.NUM_F64)
    assert a.numeric == N.NUM_I32 and b.numeric == N.NUM_F64
    assert a.alignment.f == b.alignment.f


def test_empty_alignment_cases():
    """arg-max (0,0): empty alignment, f = 0, coords ((1,1),(1,1)) — no error (pwm/mod.rs:76-81)."""
    q = np.array([0, 1, 2, 3], dtype=np.uint8)
    res = _check_single(q, -np.ones((4, 6)), 1, 1)
    assert res.alignment.f == 0.0 and len(res.alignment.numbered) == 0
    assert res.alignment.coords == ((1, 1), (1, 1))
    _check_single(q[:0], np.ones((4, 5)), 1, 1)       # empty window
    _check_single(q, np.ones((4, 0)), 1, 1)           # empty profile


def test_errors_mirror_the_reference():
    q = np.array([0, 1, 2, 3], dtype=np.uint8)
    with pytest.raises(ab.Error) as ei:
        P.PWMAligner.from_seqs(q).perform_alignment(1, 1, np.ones((5, 4)))
    assert ei.value.kind == "MatrixShapeError"
    with pytest.raises(ab.Error) as ei:
        P.PWMAligner.from_seqs(q).perform_alignment(1, 1, np.ones((4, 4)), ab.Heuristics(0, 0, np.ones(4)))
    assert ei.value.kind == "UnnecessaryArgument"
    with pytest.raises(ab.Error) as ei:
        P.HeuristicPWMAligner.from_seqs(q).perform_alignment(1, 1, np.ones((4, 4)))
    assert ei.value.kind == "MissingArgument"
    with pytest.raises(Exception):
        P.PWMAligner.from_seqs(np.array([0, 98, 1], dtype=np.uint8)).perform_alignment(1, 1, np.ones((4, 4)))
    assert P.PWMAligner.from_str_seqs("ATCG", "ignored").query.tolist() == [0, 1, 2, 3]


@pytest.mark.parametrize("numeric,d,e", [(N.NUM_AUTO, 3, 1), (N.NUM_F64, 3, 1), (N.NUM_AUTO, 2, 2)])
def test_windows_of_a_chromosome(numeric, d, e):
    seq, repeat, planted = W.dna_chromosome_with_repeats(60000, 120, 25, seed=3)
    pwm = W.pwm_from_repeat(repeat, 1.0, -1.0)
    step, win = 30, 150
    starts = np.arange(0, len(seq), step, dtype=np.uint64)
    lens = np.minimum(win, len(seq) - starts).astype(np.uint32)
    res = _check_windows(seq, starts, lens, pwm, d, e, numeric)
    assert res.f.max() > 40  # the planted copies are found


def test_strided_windows_equal_explicit_windows():
    seq, repeat, _ = W.dna_chromosome_with_repeats(20000, 60, 10, seed=4)
    pwm = W.pwm_from_repeat(repeat, 2.0, -1.0) + 0.125
    rs = P.ResidentSequence(seq)
    a = rs.align_windows(pwm, 2.5, 0.5, first=10, step=20, window=90)
    starts = np.arange(10, len(seq), 20, dtype=np.uint64)
    lens = np.minimum(90, len(seq) - starts).astype(np.uint32)
    b = rs.align_windows(pwm, 2.5, 0.5, starts=starts, lens=lens)
    rs.close()
    assert a.count == len(starts) and (a.f == b.f).all() and (a.coords == b.coords).all()
    ref = O.pwm_windows(seq, starts, lens, pwm, 2.5, 0.5, want_alignments=False)
    assert (a.f == ref["f"]).all()


def test_speculation_leftovers_and_long_windows():
    """+1/-1 profile with del 2 / ext 1 on random sequence: many H[m][x] == 0, so some windows go to
    the exact kernel; windows longer than 512 rows use several bands."""
    rng = np.random.default_rng(5)
    seq = rng.integers(0, 4, 30000).astype(np.uint8)
    pwm = P.get_random_pwm(80, rng)
    starts = np.arange(0, 29000, 97, dtype=np.uint64)
    lens = np.full(len(starts), 100, dtype=np.uint32)
    _check_windows(seq, starts, lens, pwm, 2, 1)
    starts = np.arange(0, 20000, 1500, dtype=np.uint64)
    lens = np.full(len(starts), 700, dtype=np.uint32)
    _check_windows(seq, starts, lens, pwm, 2, 1)
    _check_windows(seq, starts, lens, pwm + 0.5, 2, 1)


def test_bad_symbols_in_a_window_only_fail_that_window():
    rng = np.random.default_rng(6)
    seq = rng.integers(0, 4, 2000).astype(np.uint8)
    seq[777] = 98
    starts = np.arange(0, 1900, 50, dtype=np.uint64)
    lens = np.full(len(starts), 100, dtype=np.uint32)
    res = _check_windows(seq, starts, lens, P.get_random_pwm(60, rng), 2, 2)
    assert (res.status != 0).sum() == 2


def test_heuristic_pwm_aligner_loop():
    """heuristic/mod.rs:103-141 against the same loop run on the oracle."""
    rng = np.random.default_rng(8)
    repeat = rng.integers(0, 4, 80).astype(np.uint8)
    q = repeat.copy()
    q[rng.random(80) < 0.2] = 0
    pwm0 = W.pwm_from_repeat(repeat, 1.0, -1.0)
    fr = np.bincount(q, minlength=4) / len(q)
    h = ab.Heuristics(-1.0, 4.0 * 80, fr)
    res = P.HeuristicPWMAligner.from_seqs(q).perform_alignment(4, 4, pwm0, h)
    st, T = O.transform_matrix(pwm0, h.kd, h.r_squared, fr)
    assert st == 0
    max_f = 0.0
    while True:
        ref = O.pwm_align(q, T, 4.0, 4.0)
        if ref["f"] > max_f:
            max_f = ref["f"]
            st, T = O.transform_matrix(O.pwm_frequency_matrix(ref["numbered"], ref["q_aln"], 80), h.kd, h.r_squared, fr)
            assert st == 0
        else:
            break
    assert res.alignment.f == ref["f"]
    assert (res.alignment.numbered == ref["numbered"]).all() and (res.alignment.query == ref["q_aln"]).all()
    assert (res.matrix == T).all()
