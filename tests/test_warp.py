"""GPU parity of the warp-per-pair fill kernel (aligner_b200/csrc/kernels_warp.cuh) against the
oracle: every rows-per-lane instantiation, several bands, the stateful-penalty speculation
(including pairs that never settle and fall through to the exact kernel), INT32 and FP64."""
import numpy as np
import pytest

import oracle_lib as O

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W
from test_gpu_parity import _check_batch, _check_pair

pytestmark = pytest.mark.gpu


def _ragged(rng, count, lo, hi, V, related=True):
    qs, ts = [], []
    for _ in range(count):
        n = int(rng.integers(lo, hi + 1))
        q = rng.integers(0, V, n).astype(np.uint8)
        if related:
            t = q.copy()
            mut = rng.random(n) < 0.25
            t[mut] = rng.integers(0, V, int(mut.sum()))
            cut = int(rng.integers(0, max(1, n // 8)))
            t = t[cut:] if cut < n - 1 else t
        else:
            t = rng.integers(0, V, int(rng.integers(lo, hi + 1))).astype(np.uint8)
        qs.append(q)
        ts.append(t)
    return ab.PackedPairs.from_lists(qs, ts)


@pytest.mark.parametrize("hi", [40, 90, 100, 150, 180, 220, 250, 280, 310, 340, 380, 410, 440, 470, 512])
def test_every_rows_per_lane_int(hi):
    """max target length picks R = 2, 3, ..., 16 (one band)."""
    rng = np.random.default_rng(hi)
    pairs = _ragged(rng, 96, max(24, hi - 60), hi, 24)
    for mode, d, e in [("local", 11, 2), ("local", 5, 5), ("global", 11, 1)]:
        res = _check_batch(mode, pairs, d, e, ab.get_blosum62(), numeric=N.NUM_I32)
        assert (res.status == 0).all()


@pytest.mark.parametrize("force_mode2", [False, True])
@pytest.mark.parametrize("hi", [40, 100, 180, 250, 310, 380, 440, 512])
def test_int32_static_penalty_both_modes(hi, force_mode2, monkeypatch):
    """INT32 local alignment with del == ext through the compile-time static-penalty instantiation
    (MODE 1, the default again) and through the MODE 2 code (ALNR_WARP_INT_MODE2=1): round 1 saw
    wrong MODE 1 scores on an intermediate build and routed around it; both must be bit-exact."""
    if force_mode2:
        monkeypatch.setenv("ALNR_WARP_INT_MODE2", "1")
    else:
        monkeypatch.delenv("ALNR_WARP_INT_MODE2", raising=False)
    rng = np.random.default_rng(7000 + hi)
    pairs = _ragged(rng, 64, max(24, hi - 60), hi, 24)
    for d in (5, 11):
        res = _check_batch("local", pairs, d, d, ab.get_blosum62(), numeric=N.NUM_I32)
        assert (res.status == 0).all()
    dna = _ragged(rng, 64, max(24, hi - 60), hi, 4, related=False)
    _check_batch("local", dna, 1, 1, ab.dna_matrix(1, -1), numeric=N.NUM_I32)  # many ties and zeros


@pytest.mark.parametrize("hi", [60, 90, 150, 200, 220, 280, 330, 410, 470, 520])
def test_every_rows_per_lane_f64(hi):
    rng = np.random.default_rng(1000 + hi)
    pairs = _ragged(rng, 48, max(24, hi - 60), hi, 24)
    M = ab.get_blosum62() + rng.normal(0, 0.37, (24, 24))
    for mode, d, e in [("local", 4.25, 4.25), ("local", 6.5, 1.125), ("global", 3.3, 0.7)]:
        _check_batch(mode, pairs, d, e, M)


def test_multi_band_single_pairs():
    """targets longer than 32*16 rows: several bands through the row buffer."""
    rng = np.random.default_rng(31)
    for n, m in [(700, 1500), (1300, 530), (90, 2049)]:
        q = rng.integers(0, 4, n).astype(np.uint8)
        t = rng.integers(0, 4, m).astype(np.uint8)
        k = min(n, m)
        t[:k] = np.where(rng.random(k) < 0.8, q[:k], t[:k])
        for mode, d, e in [("local", 10, 2), ("global", 8, 8), ("local", 3, 3)]:
            _check_pair(mode, q, t, d, e, ab.dna_matrix(5, -4), ab.DNA)
        _check_pair("local", q, t, 2.5, 0.75, ab.dna_matrix(5, -4) + 0.1, ab.DNA)


def test_multi_band_batch():
    rng = np.random.default_rng(32)
    pairs = _ragged(rng, 40, 500, 1100, 4)
    _check_batch("local", pairs, 10, 2, ab.dna_matrix(5, -4))
    _check_batch("global", pairs, 10, 2, ab.dna_matrix(5, -4))


def test_speculation_never_settles_goes_to_exact_kernel():
    """+1/-1, del=2, ext=1 on unrelated sequences: H[m][x] == 0 in many columns, so some pairs need
    more passes than the warp kernel makes and are finished by the exact column-major kernel."""
    rng = np.random.default_rng(33)
    pairs = _ragged(rng, 400, 30, 120, 4, related=False)
    _check_batch("local", pairs, 2, 1, ab.dna_matrix(1, -1), numeric=N.NUM_I32)
    _check_batch("local", pairs, 2, 1, ab.dna_matrix(1, -1), numeric=N.NUM_F64)
    # single pair with full matrices through the same path
    q, t = pairs.pair(7)
    _check_pair("local", q, t, 2, 1, ab.dna_matrix(1, -1), ab.DNA)


def test_protein_pairs_config5_shape_int_and_f64():
    pairs = W.protein_pairs(0, 128, 270, 330)
    a = _check_batch("local", pairs, 5, 5, ab.get_blosum62(), numeric=N.NUM_I32)
    b = _check_batch("local", pairs, 5, 5, ab.get_blosum62(), numeric=N.NUM_F64)
    assert (a.f == b.f).all()


def test_large_ragged_batch_uses_sorted_chunks():
    rng = np.random.default_rng(34)
    pairs = _ragged(rng, 6000, 24, 90, 4)
    _check_batch("local", pairs, 7, 3, ab.dna_matrix(3, -2))
