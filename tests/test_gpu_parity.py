"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same
inputs.  Bar: bit-exact scores, coordinates, aligned strings and (single-pair calls) the
full score and direction matrices."""
import numpy as np
import pytest

import oracle_lib as O
from conftest import load_golden, sha16

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W

pytestmark = pytest.mark.gpu

MODE = {"global": N.GLOBAL, "local": N.LOCAL}
CLS = {"global": ab.SimpleGlobalAligner, "local": ab.SimpleLocalAligner}


def _check_pair(mode, q, t, d, e, M, symbols, numeric=N.NUM_AUTO):
    ref = O.align_pair(mode, q, t, float(d), float(e), M)
    al = CLS[mode].from_seqs(q, t, symbols=symbols)
    if ref["status"] != 0:
        with pytest.raises((ab.AlignerError, ab.Error)) as ei:
            al.perform_alignment(d, e, M, numeric=numeric)
        assert getattr(ei.value, "status", None) == O.STATUS_TO_ALNR[ref["status"]]
        return None
    r = al.perform_alignment(d, e, M, numeric=numeric)
    assert (r.alignment_matrix == ref["H"]).all()
    assert (r.direction_matrix == ref["D"]).all()
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()
    assert r.alignment.coords == ref["coords"]
    assert r.alignment.f == ref["f"] and r.corner_score == ref["corner"]
    return r


def test_config1_book_global_legacy_golden():
    """BASELINE config 1 + the reference's golden vector src/tests/test_alignment.rs:14-39."""
    g = load_golden("legacy_global.json")
    M = np.zeros((24, 24))
    for tl, row in g["blosum50_sub"].items():
        for ql, v in row.items():
            M[ab.Protein.match_with_char(tl), ab.Protein.match_with_char(ql)] = v
    al = ab.SimpleGlobalAligner.from_str_seqs(g["query"], g["target"])
    r = al.perform_alignment(8, 8, M)
    assert (r.alignment_matrix == np.array(g["H"], dtype=np.float64)).all()
    assert (r.direction_matrix == np.array(g["D"], dtype=np.uint8)).all()
    assert r.alignment.strings() == ("HEAGAWGHE_EE", "_PA__W_HEAEE")
    assert r.alignment.f == 0.0 and r.alignment.coords == ((1, 10), (1, 7))


@pytest.mark.parametrize("mode,d,e", [("global", 8, 8), ("global", 11, 2), ("local", 11, 2), ("local", 8, 8)])
def test_config1_book_blosum62(examples, mode, d, e):
    q, t = examples["book"]
    r = _check_pair(mode, q, t, d, e, ab.get_blosum62(), ab.Protein)
    assert r.numeric in (N.NUM_I32, N.NUM_I16X2)


@pytest.mark.parametrize("mode,d,e", [("local", 11, 2), ("local", 11, 1), ("local", 11, 11), ("global", 11, 2)])
def test_config2_protein_fasta(examples, mode, d, e):
    q, t = examples["protein"]
    r = _check_pair(mode, q, t, d, e, ab.get_blosum62(), ab.Protein)
    if (mode, d, e) == ("local", 11, 2):
        assert r.alignment.f == 1554.0 and len(r.alignment.query) == 342
        assert sha16(r.direction_matrix) == "21a6e7676bbc9088"


@pytest.mark.parametrize("mode,d,e", [("global", 8, 8), ("local", 10, 2)])
def test_config4_gene_file(examples, mode, d, e):
    q, t = examples["gene"]
    _check_pair(mode, q, t, d, e, ab.dna_matrix(5, -4), ab.DNA)


@pytest.mark.parametrize("numeric", [N.NUM_I32, N.NUM_F64])
def test_forced_domains_agree(examples, numeric):
    q, t = examples["book"]
    for mode in ("global", "local"):
        _check_pair(mode, q, t, 11, 2, ab.get_blosum62(), ab.Protein, numeric=numeric)


def test_fp64_non_integral_matrix():
    rng = np.random.default_rng(5)
    for it in range(40):
        V = 4 if it % 2 else 24
        n, m = int(rng.integers(1, 40)), int(rng.integers(1, 40))
        q = rng.integers(0, V, n).astype(np.uint8)
        t = rng.integers(0, V, m).astype(np.uint8)
        M = rng.normal(0, 2, (V, V))
        d, e = float(rng.uniform(0, 6)), float(rng.uniform(0, 3))
        for mode in ("global", "local"):
            r = _check_pair(mode, q, t, d, e, M, ab.DNA if V == 4 else ab.Protein)
            if r is not None:
                assert r.numeric == N.NUM_F64


def test_status_cases():
    M = ab.dna_matrix()
    a = np.array([0, 1, 2], dtype=np.uint8)
    for mode, q, t, mat in [("global", a[:0], a, M), ("local", a, a[:0], M),
                            ("local", np.array([98], dtype=np.uint8), a, M), ("local", a, a, -np.ones((4, 4)))]:
        _check_pair(mode, q, t, 1, 1, mat, ab.DNA)


def _check_batch(mode, pairs, d, e, M, numeric=N.NUM_AUTO, threads=8):
    ref = O.align_batch(mode, pairs, float(d), float(e), M, threads=threads)
    res = ab.align_batch(MODE[mode], pairs, M, d, e, numeric=numeric)
    want_status = np.array([O.STATUS_TO_ALNR[int(s)] for s in ref["status"]], dtype=np.int32)
    assert (res.status == want_status).all()
    ok = want_status == 0
    assert (res.f[ok] == ref["f"][ok]).all()
    assert (res.corner[ok] == ref["corner"][ok]).all()
    assert (res.coords[ok] == ref["coords"][ok]).all()
    assert (res.aln_len[ok] == ref["aln_len"][ok]).all()
    for i in np.nonzero(ok)[0]:
        qa, ta = res.alignment(i)
        L = int(ref["aln_len"][i])
        assert (qa == ref["q_aln"][i, :L]).all() and (ta == ref["t_aln"][i, :L]).all(), i
    return res


@pytest.mark.parametrize("mode,d,e", [("local", 6, 6), ("local", 10, 2), ("global", 6, 6), ("global", 10, 2)])
def test_config3_batch_subsample(mode, d, e):
    """BASELINE config 3 on a subsample: 150 bp DNA read pairs, +5/-4."""
    pairs = W.dna_read_batch(0, 3000, n=150)
    _check_batch(mode, pairs, d, e, ab.dna_matrix(5, -4), numeric=N.NUM_I16X2)  # the kernel of the full-size run
    _check_batch(mode, pairs, d, e, ab.dna_matrix(5, -4))                        # what AUTO picks for 3000 pairs


def test_ragged_batch_with_bad_pairs():
    rng = np.random.default_rng(9)
    qs, ts = [], []
    for i in range(500):
        n, m = int(rng.integers(0 if i % 50 == 0 else 1, 200)), int(rng.integers(1, 200))
        q = rng.integers(0, 4, n).astype(np.uint8)
        t = rng.integers(0, 4, m).astype(np.uint8)
        if i % 77 == 5 and n:
            q[n // 2] = 98
        qs.append(q)
        ts.append(t)
    pairs = ab.PackedPairs.from_lists(qs, ts)
    for mode in ("local", "global"):
        _check_batch(mode, pairs, 7, 3, ab.dna_matrix(3, -2))


def test_protein_batch_i32():
    pairs = W.protein_pairs(0, 64, 40, 90)
    _check_batch("local", pairs, 11, 2, ab.get_blosum62())
    _check_batch("global", pairs, 11, 1, ab.get_blosum62())


def test_device_batch_rerun_is_idempotent():
    pairs = W.dna_read_batch(100, 512, n=150)
    db = ab.DeviceBatch(N.LOCAL, pairs, ab.dna_matrix(), 10, 2)
    db.run(); db.sync()
    a = db.download()
    ms, _ = db.run_timed()
    b = db.download()
    assert ms > 0
    for k in ("f", "coords", "aln_len", "aln_off", "q_aln", "t_aln"):
        assert (getattr(a, k) == getattr(b, k)).all()
    db.close()


# ---- the packed 16-bit strip kernel (two pairs per thread) -------------------------------
def _rand_uniform_pairs(rng, count, n, m, related=True):
    q = rng.integers(0, 4, (count, n)).astype(np.uint8)
    if related and n == m:
        t = np.where(rng.random((count, n)) < 0.12, rng.integers(0, 4, (count, n)), q).astype(np.uint8)
    else:
        t = rng.integers(0, 4, (count, m)).astype(np.uint8)
    return ab.PackedPairs.uniform(q, t)


@pytest.mark.parametrize("n,m,count", [(150, 150, 640), (150, 150, 33), (37, 61, 200), (64, 16, 130),
                                       (1, 1, 70), (5, 150, 64), (151, 149, 257), (300, 31, 65)])
@pytest.mark.parametrize("mode,d,e", [("local", 10, 2), ("local", 6, 6), ("global", 10, 2), ("global", 4, 4)])
def test_strip_kernel_shapes(n, m, count, mode, d, e):
    rng = np.random.default_rng(n * 1000 + m + count)
    pairs = _rand_uniform_pairs(rng, count, n, m)
    _check_batch(mode, pairs, d, e, ab.dna_matrix(5, -4), numeric=N.NUM_I16X2)


def test_strip_kernel_many_speculation_failures():
    """+1/-1 scores with del=2, ext=1 on unrelated pairs: H[m][x] == 0 is frequent, so the
    speculative pass fails for most pairs and the re-run passes / exact leftovers do the work."""
    rng = np.random.default_rng(123)
    pairs = _rand_uniform_pairs(rng, 1500, 60, 45, related=False)
    M = ab.dna_matrix(1, -1)
    ref = O.align_batch("local", pairs, 2.0, 1.0, M, threads=8)
    db = ab.DeviceBatch(N.LOCAL, pairs, M, 2, 1, numeric=N.NUM_I16X2)
    db.run(); db.sync()
    res = db.download()
    assert db.numeric == N.NUM_I16X2 and db.rerun_count > 100
    ok = ref["status"] == 0
    assert ((res.status == 0) == ok).all()
    assert (res.f[ok] == ref["f"][ok]).all() and (res.coords[ok] == ref["coords"][ok]).all()
    for i in np.nonzero(ok)[0]:
        qa, ta = res.alignment(i)
        L = int(ref["aln_len"][i])
        assert (qa == ref["q_aln"][i, :L]).all() and (ta == ref["t_aln"][i, :L]).all(), i
    db.close()


def test_strip_matches_exact_domains_on_asymmetric_matrix():
    rng = np.random.default_rng(77)
    pairs = _rand_uniform_pairs(rng, 300, 90, 90)
    M = rng.integers(-6, 8, (4, 4)).astype(np.float64)
    for mode, d, e in [("local", 9, 3), ("global", 7, 2)]:
        a = _check_batch(mode, pairs, d, e, M, numeric=N.NUM_I16X2)
        b = _check_batch(mode, pairs, d, e, M, numeric=N.NUM_I32)
        c = _check_batch(mode, pairs, d, e, M, numeric=N.NUM_F64)
        assert (a.f == b.f).all() and (b.f == c.f).all()


def test_auto_selects_strip_for_config3():
    """Packed strip kernel from 16384 uniform DNA pairs on; smaller batches run one warp per pair."""
    pairs = W.dna_read_batch(0, 16384, n=150)
    db = ab.DeviceBatch(N.LOCAL, pairs, ab.dna_matrix(), 10, 2)
    assert db.numeric == N.NUM_I16X2
    db.close()
    db = ab.DeviceBatch(N.LOCAL, W.dna_read_batch(0, 256, n=150), ab.dna_matrix(), 10, 2)
    assert db.numeric == N.NUM_I32
    db.close()
