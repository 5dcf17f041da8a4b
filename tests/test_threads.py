"""The reference's callers invoke perform_alignment concurrently from plain OS threads with
cloned inputs (statistics/mod.rs:255-286, latent-repeat-search/engine/calc.rs:47-69): the C ABI
must be callable from many host threads at once ("Threading")."""
import threading

import numpy as np
import pytest

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _jobs():
    rng = np.random.default_rng(5)
    M62 = ab.get_blosum62()
    Md = ab.dna_matrix(5, -4)
    prot = W.protein_pairs(3, 6, 200, 340)
    reads = W.dna_read_batch(1, 512, n=150)
    lq, lt = W.dna_long_pair(2500, seed=9)
    jobs = []
    for i in range(6):
        q, t = prot.pair(i)
        jobs.append(lambda q=q, t=t: _pair(ab.SimpleLocalAligner.from_seqs(q, t).perform_alignment(11, 2, M62)))
        jobs.append(lambda q=q, t=t: _pair(ab.SimpleGlobalAligner.from_seqs(q, t).perform_alignment(8, 8, M62)))
    jobs.append(lambda: _batch(ab.align_batch(N.LOCAL, reads, Md, 10, 2)))
    jobs.append(lambda: _batch(ab.align_batch(N.GLOBAL, reads, Md, 6, 6)))
    jobs.append(lambda: _pair(ab.SimpleGlobalAligner.from_seqs(lq, lt, symbols=ab.DNA).perform_alignment(
        8, 8, Md, want_matrices=False)))
    jobs.append(lambda: _pair(ab.SimpleLocalAligner.from_seqs(lq, lt, symbols=ab.DNA).perform_alignment(
        6, 6, Md, want_matrices=False)))
    sc = ab.score_batch  # score-only batch (statistics shuffle path)
    jobs.append(lambda: (sc(N.LOCAL, reads, Md, 10, 2).f.tobytes(),))
    hp = W.protein_pairs(8, 12, 120, 200)
    mats = W.candidate_matrices(6)
    freqs = W.pair_frequencies(hp, 24)

    def heur():  # the matrix-search loop (heuristic/mod.rs:36-78) over a 12 x 6 grid
        r = ab.heuristic_batch(hp, mats, 5, 5, -1.0, 576.0, freqs, want_matrices=False)
        return (r.f.tobytes(), r.passes.tobytes(), r.status.tobytes())
    jobs.append(heur)
    Mf = M62 + rng.normal(0, 0.25, (24, 24))
    jobs.append(lambda: _batch(ab.align_batch(N.LOCAL, prot, Mf, 4.5, 1.25)))
    return jobs


def _pair(r):
    return (r.alignment.f, r.corner_score, r.alignment.coords, r.alignment.query.tobytes(), r.alignment.target.tobytes())


def _batch(res):
    return (res.f.tobytes(), res.corner.tobytes(), res.coords.tobytes(), res.aln_len.tobytes(),
            b"".join(res.alignment(i)[0].tobytes() + res.alignment(i)[1].tobytes() for i in range(0, len(res.f), 37)))


def test_concurrent_calls_from_host_threads():
    jobs = _jobs()
    expected = [j() for j in jobs]  # one after the other
    n_threads, rounds = 8, 3
    results = [[None] * len(jobs) for _ in range(n_threads)]
    errors = []
    start = threading.Barrier(n_threads)

    def worker(k):
        try:
            start.wait()
            for r in range(rounds):
                for idx in np.random.default_rng(100 * k + r).permutation(len(jobs)):
                    results[k][idx] = jobs[idx]()
        except Exception as e:  # noqa: BLE001
            errors.append((k, repr(e)))

    ths = [threading.Thread(target=worker, args=(k,)) for k in range(n_threads)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errors, errors
    for k in range(n_threads):
        for idx in range(len(jobs)):
            assert results[k][idx] == expected[idx], (k, idx)
