"""Multi-GPU partitioning behind the C ABI (alnr_align_batch_multi / alnr_matrix_search_multi,
aligner_b200/csrc/multi.cu): any number of shards must give the bytes of the single-device call.
Shard k runs on device k % device_count, so the sharding logic is exercised on a one-GPU box too."""
import numpy as np
import pytest

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W
from test_compact import ragged_dna

pytestmark = pytest.mark.gpu
M = ab.dna_matrix(5, -4)


def _same(a, b, pairs):
    assert (a.f == b.f).all() and (a.corner == b.corner).all() and (a.status == b.status).all()
    assert (a.coords == b.coords).all() and (a.aln_len == b.aln_len).all()
    for i in range(pairs.count):
        qa, ta = a.alignment(i)
        qb, tb = b.alignment(i)
        assert (qa == qb).all() and (ta == tb).all(), i


@pytest.mark.parametrize("n_gpus", [2, 3, 0])
def test_align_batch_multi_uniform_strings(n_gpus):
    pairs = W.dna_read_batch(0, 40000, n=150)
    ref = ab.align_batch(N.LOCAL, pairs, M, 10, 2)
    res = ab.align_batch(N.LOCAL, pairs, M, 10, 2, n_gpus=n_gpus)
    _same(res, ref, pairs)


@pytest.mark.parametrize("packed", [False, True])
def test_align_batch_multi_compact(packed):
    pairs = W.dna_read_batch(500, 40000, n=150)
    src = pairs.pack2bit() if packed else pairs
    ref = ab.align_batch(N.LOCAL, pairs, M, 10, 2)
    cr = ab.align_batch(N.LOCAL, src, M, 10, 2, compact=True, n_gpus=3)
    assert (cr.f == ref.f).all() and (cr.aln_len == ref.aln_len).all() and (cr.coords == ref.coords).all()
    ex = cr.expand(threads=4)
    _same(ex, ref, pairs)


def test_align_batch_multi_ragged_global():
    pairs = ragged_dna(900, 10, 300, seed=21)
    ref = ab.align_batch(N.GLOBAL, pairs, M, 8, 8)
    res = ab.align_batch(N.GLOBAL, pairs, M, 8, 8, n_gpus=4)
    _same(res, ref, pairs)
    cr = ab.align_batch(N.GLOBAL, pairs.pack2bit(), M, 8, 8, compact=True, n_gpus=2)
    _same(cr.expand(), ref, pairs)


def test_matrix_search_multi_bit_identical():
    pairs = W.protein_pairs(0, 700, 60, 90)
    mats = W.candidate_matrices(6)
    fr = W.pair_frequencies(pairs, 24)
    ref = ab.matrix_search(pairs, mats, 5, 5, -1.0, 0.0, fr)
    for n in (2, 3):
        got = ab.matrix_search(pairs, mats, 5, 5, -1.0, 0.0, fr, n_gpus=n)
        assert got[0].tobytes() == ref[0].tobytes()  # fitness bits
        assert (got[1] == ref[1]).all() and (got[2] == ref[2]).all()
        assert (got[3] == ref[3]).all() and (got[4] == ref[4]).all()
