"""Heuristic matrix-search path: transform_matrix (host instance, CPU test) and the GPU loop
against the oracle (oracle_heur.cpp / oracle_matrix.cpp; parity unpinned)."""
import numpy as np
import pytest

import oracle_lib as O

import aligner_b200 as ab
from aligner_b200 import workloads as W


def _cases(rng, count):
    for it in range(count):
        V = 24 if it % 2 else 4
        M = rng.integers(-1, 2, (V, V)).astype(np.float64) if it % 3 else rng.normal(0, 1, (V, V))
        fr = rng.uniform(0.5, 1.5, V)
        fr /= fr.sum()
        yield M, -0.9 + 0.01 * it, float(V * V), fr


def test_transform_matrix_host_matches_oracle_bitwise():
    rng = np.random.default_rng(21)
    n_ok = 0
    for M, kd, r2, fr in _cases(rng, 60):
        st, T = O.transform_matrix(M, kd, r2, fr)
        if st:
            with pytest.raises(ab.Error):
                ab.transform_matrix(M, kd, r2, fr)
            continue
        assert (ab.transform_matrix(M, kd, r2, fr) == T).all()
        n_ok += 1
    assert n_ok > 30
    # non-square (PWM-shaped) matrices too: 4 x L
    M = rng.integers(-1, 2, (4, 37)).astype(np.float64)
    fr = np.array([0.3, 0.2, 0.25, 0.25])
    st, T = O.transform_matrix(M, 0.0, 210.0, fr)
    assert st == 0 and (ab.transform_matrix(M, 0.0, 210.0, fr) == T).all()


def test_heuristic_requires_heuristics_argument():
    a = ab.HeuristicAligner.from_str_seqs("HEAGAWGHEE", "PAWHEAE")
    with pytest.raises(ab.Error) as ei:
        a.perform_alignment(5, 5, ab.get_blosum62())
    assert ei.value.kind == "MissingArgument"  # heuristic/mod.rs:43-46


def _oracle_items(pairs, mats, d, e, kd, r2, freqs, pi, mi):
    out = []
    for p, k in zip(pi, mi):
        q, t = pairs.pair(int(p))
        out.append(O.heuristic(q, t, float(d), float(e), mats[int(k)], kd, r2, freqs[int(p)]))
    return out


@pytest.mark.gpu
def test_heuristic_single_pair_matches_oracle(examples):
    q, t = examples["protein"]
    q, t = q[:120], t[:110]
    M = W.candidate_matrices(3)[1]
    fr = np.bincount(np.concatenate([q, t]), minlength=24).astype(np.float64)
    fr /= fr.sum()
    ref = O.heuristic(q, t, 5.0, 5.0, M, -1.0, 0.0, fr)
    assert ref["status"] == 0 and ref["passes"] >= 2
    r = ab.HeuristicAligner.from_seqs(q, t).perform_alignment(5, 5, M, ab.Heuristics(-1.0, 0.0, fr))
    assert r.alignment.f == ref["f"] and r.passes == ref["passes"]
    assert (r.matrix == ref["matrix"]).all()
    assert (r.alignment.query == ref["q_aln"]).all() and (r.alignment.target == ref["t_aln"]).all()
    assert r.alignment.coords == ref["coords"]


@pytest.mark.gpu
def test_heuristic_grid_matches_oracle_config5_subgrid():
    """BASELINE config 5 on the sub-grid SURVEY.md 8d names: 10 protein pairs of ~300 aa x 50
    candidate matrices (500 items, each the oracle's full heuristic loop)."""
    pairs = W.protein_pairs(0, 10, 270, 330)
    mats = W.candidate_matrices(50)
    freqs = W.pair_frequencies(pairs, 24)
    fitness, best_k, best_f, grid, sgrid = ab.matrix_search(pairs, mats, 5, 5, -1.0, 576.0, freqs)
    P, K = pairs.count, len(mats)
    ref_f = np.zeros((P, K))
    ref_s = np.zeros((P, K), dtype=np.int32)
    for p in range(P):
        q, t = pairs.pair(p)
        for k in range(K):
            r = O.heuristic(q, t, 5.0, 5.0, mats[k], -1.0, 576.0, freqs[p])
            ref_s[p, k] = O.STATUS_TO_ALNR[r["status"]]
            ref_f[p, k] = r["f"] if r["status"] == 0 else 0.0
    assert (sgrid == ref_s).all()
    ok = ref_s == 0
    assert ok.sum() > P * K // 2
    assert (grid[ok] == ref_f[ok]).all()
    # fitness: per matrix, f added in pair order
    for k in range(K):
        acc = 0.0
        for p in range(P):
            if ok[p, k]:
                acc = acc + ref_f[p, k]
        assert fitness[k] == acc
    for p in range(P):
        if ok[p].any():
            masked = np.where(ok[p], ref_f[p], -np.inf)
            assert best_k[p] == int(np.argmax(masked)) and best_f[p] == masked.max()


@pytest.mark.gpu
def test_heuristic_items_list_dna_with_matrices():
    rng = np.random.default_rng(4)
    pairs = W.dna_read_batch(0, 40, n=50)
    mats = rng.integers(-1, 2, (6, 4, 4)).astype(np.float64)
    freqs = W.pair_frequencies(pairs, 4)
    pi = rng.integers(0, 40, 100).astype(np.uint32)
    mi = rng.integers(0, 6, 100).astype(np.uint32)
    res = ab.heuristic_batch(pairs, mats, 3.0, 3.0, -0.5, 0.0, freqs, pi, mi)
    ref = _oracle_items(pairs, mats, 3.0, 3.0, -0.5, 0.0, freqs, pi, mi)
    n_ok = 0
    for i, r in enumerate(ref):
        assert res.status[i] == O.STATUS_TO_ALNR[r["status"]], i
        if r["status"] == 0:
            n_ok += 1
            assert res.f[i] == r["f"] and res.passes[i] == r["passes"]
            assert (res.matrices[i] == r["matrix"]).all()
            assert tuple(int(v) for v in res.coords[i]) == r["coords"][0] + r["coords"][1]
    assert n_ok > 20
