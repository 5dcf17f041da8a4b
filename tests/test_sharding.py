"""Multi-GPU layer (aligner_b200/sharding.py): shard boundaries, the fixed-order fitness
reduction (bit-identical for every world size) over gloo with world_size 2 and 3 on CPU, and —
on the GPU box — the sharded search / batch with two ranks sharing cuda:0."""
import os
import pickle
import socket
import subprocess
import sys

import numpy as np
import pytest

import aligner_b200 as ab
from aligner_b200 import sharding as S
from aligner_b200 import workloads as W

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _launch(world, mode, tmp_path, *args):
    out = str(tmp_path / ("%s_%d.pkl" % (mode, world)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(HERE, "_dist_worker.py"), mode, out] + [str(a) for a in args]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    with open(out, "rb") as fh:
        return pickle.load(fh)


def test_shard_bounds_cover_and_balance():
    rng = np.random.default_rng(3)
    for world in (1, 2, 3, 4, 8):
        for count in (0, 1, 5, 255, 256, 257, 5000):
            w = rng.integers(1, 100, count).astype(np.float64)
            for align in (1, 64, S.BLOCK):
                b = S.shard_bounds(w, world, align)
                assert len(b) == world + 1 and b[0] == 0 and b[-1] == count
                assert all(x <= y for x, y in zip(b, b[1:]))
                assert all(x % align == 0 for x in b[1:-1])
            if count == 5000:
                b = S.shard_bounds(w, world, 1)
                loads = [w[x:y].sum() for x, y in zip(b, b[1:])]
                assert max(loads) - min(loads) <= 2 * w.max()


def test_slice_pairs_rebases_offsets():
    pairs = W.protein_pairs(0, 20, 10, 30)
    sub = S.slice_pairs(pairs, 5, 12)
    assert sub.count == 7 and sub.q_off[0] == 0 and sub.t_off[0] == 0
    for i in range(7):
        q, t = sub.pair(i)
        q0, t0 = pairs.pair(5 + i)
        assert (q == q0).all() and (t == t0).all()


def test_fitness_blocks_fixed_order():
    from _dist_worker import synthetic_grid
    f, st = synthetic_grid(1000, 5)
    blocks = S.fitness_block_sums(f, st)
    assert blocks.shape == (4, 5)
    # restated: sequential sums inside a block, blocks added in order
    want = np.zeros(5)
    for b in range(4):
        acc = np.zeros(5)
        for i in range(b * 256, min(1000, (b + 1) * 256)):
            acc = np.where(st[i] == 0, acc + f[i], acc)
        assert (blocks[b] == acc).all()
        want = want + acc
    assert (S.fitness_from_blocks(blocks) == want).all()


@pytest.mark.parametrize("world", [2, 3])
def test_fitness_reduction_is_world_size_invariant_gloo(world, tmp_path):
    from _dist_worker import synthetic_grid
    P, K = 2000, 9
    f, st = synthetic_grid(P, K)
    single = S.fitness_from_blocks(S.fitness_block_sums(f, st))
    got = _launch(world, "fitness", tmp_path, P, K)
    assert len(got) == world
    ranges = [g["range"] for g in got]
    assert ranges[0][0] == 0 and ranges[-1][1] == P
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    for g in got:  # every rank holds the same bits as the single-process sum
        assert g["fitness"].tobytes() == single.tobytes()


@pytest.mark.gpu
def test_matrix_search_sharded_matches_single_gpu(tmp_path):
    P, K = 600, 6
    pairs = W.protein_pairs(0, P, 40, 70)
    mats = W.candidate_matrices(K)
    freqs = W.pair_frequencies(pairs, 24)
    fitness, best_k, best_f, grid, sgrid = ab.matrix_search(pairs, mats, 5, 5, -1.0, 576.0, freqs)
    fit1, (lo, hi), bk, bf, f, st = S.matrix_search_sharded(pairs, mats, 5, 5, -1.0, 576.0, freqs)
    assert (lo, hi) == (0, P) and fit1.tobytes() == fitness.tobytes()
    assert (bk == best_k).all() and (f == grid).all()
    got = _launch(2, "search", tmp_path, P, K)  # two ranks, both on cuda:0, gloo for the reduction
    assert got[0]["range"][1] == got[1]["range"][0] and got[0]["range"][1] % S.BLOCK == 0
    for g in got:
        assert g["fitness"].tobytes() == fitness.tobytes()
        lo, hi = g["range"]
        assert (g["f"] == grid[lo:hi]).all() and (g["best_k"] == best_k[lo:hi]).all()


@pytest.mark.gpu
def test_align_batch_sharded_two_ranks(tmp_path):
    P = 3000
    pairs = W.dna_read_batch(0, P, n=150)
    ref = ab.align_batch(ab.LOCAL, pairs, ab.dna_matrix(5, -4), 10, 2)
    got = _launch(2, "batch", tmp_path, P)[0]
    assert len(got) == 2 and got[0][0] == 0 and got[0][1] == got[1][0] and got[1][1] == P
    for lo, hi, f, coords, aln_len, compact_same, strings in got:
        assert (f == ref.f[lo:hi]).all() and (coords == ref.coords[lo:hi]).all()
        assert (aln_len == ref.aln_len[lo:hi]).all()
        assert compact_same  # the compact gather expands to the gathered strings
        for i, (qa, ta) in zip((0, (hi - lo) // 2, hi - lo - 1), strings):
            qr, tr = ref.alignment(lo + i)
            assert (qa == qr).all() and (ta == tr).all()


@pytest.mark.parametrize("world", [2, 3])
def test_gather_arrays_without_pickling_gloo(world, tmp_path):
    """sharding._gather_arrays: ragged per-rank arrays (different lengths, an empty one, 2-D) arrive
    on rank 0 bit for bit; the other ranks receive nothing."""
    got = _launch(world, "gather", tmp_path)
    assert all(g["got"] is None for g in got[1:])
    for r in range(world):
        for sent, recv in zip(got[r]["sent"], got[0]["got"][r]):
            assert recv.dtype == sent.dtype and recv.shape == sent.shape and (recv == sent).all()


@pytest.mark.parametrize("world", [2, 3])
def test_window_ranges_tile_the_scan_gloo(world, tmp_path):
    got = _launch(world, "windows", tmp_path, 1001)
    assert got[0][0] == 0 and got[-1][1] == 1001
    assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
    assert max(b - a for a, b in got) - min(b - a for a, b in got) <= 1


@pytest.mark.gpu
def test_pwm_scan_and_score_batch_sharded_two_ranks(tmp_path):
    from aligner_b200 import pwm as P
    seq, repeat, _ = W.dna_chromosome_with_repeats(30000, 80, 12, seed=2)
    pwm = W.pwm_from_repeat(repeat, 1.0, -1.0) + 0.25
    rs = P.ResidentSequence(seq)
    ref = rs.align_windows(pwm, 3.0, 1.0, first=5, step=20, window=100).f
    rs.close()
    single = S.pwm_scan_sharded(seq, pwm, 3.0, 1.0, first=5, step=20, window=100)
    assert (single == ref).all()
    for f in _launch(2, "pwm", tmp_path, 30000):  # every rank holds the whole scan
        assert (f == ref).all()
    pairs = W.protein_pairs(0, 500, 40, 90)
    want = ab.score_batch(ab.LOCAL, pairs, ab.get_blosum62(), 11, 2)
    for f, st in _launch(2, "scores", tmp_path, 500):
        assert (f == want.f).all() and (st == want.status).all()
