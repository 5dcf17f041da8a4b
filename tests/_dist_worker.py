"""Worker for tests/test_sharding.py: one rank of a world_size-N gloo job (launched by
torch.distributed.run on 127.0.0.1).  mode "fitness": synthetic f grid, host math only (CPU test);
mode "search" / "batch": the real sharded calls, every rank on cuda:0 (GPU test on a 1-GPU box)."""
import os
import pickle
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch.distributed as dist  # noqa: E402

import aligner_b200 as ab  # noqa: E402
from aligner_b200 import sharding as S  # noqa: E402
from aligner_b200 import workloads as W  # noqa: E402


def synthetic_grid(P, K, seed=11):
    rng = np.random.default_rng(seed)
    f = rng.normal(50.0, 17.0, (P, K)) * rng.choice([1e-3, 1.0, 1e6], (P, K))
    st = (rng.random((P, K)) < 0.03).astype(np.int32) * 11
    return f, st


def main():
    mode, out_path = sys.argv[1], sys.argv[2]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    result = None
    if mode == "fitness":
        P, K = int(sys.argv[3]), int(sys.argv[4])
        f, st = synthetic_grid(P, K)
        w = np.arange(P, dtype=np.float64) % 7 + 1
        lo, hi = S.shard_range(P, rank, world, w, align=S.BLOCK)
        blocks = S.fitness_block_sums(f[lo:hi], st[lo:hi]) if hi > lo else np.zeros((0, K))
        fit = S.reduce_fitness(blocks, lo // S.BLOCK, (P + S.BLOCK - 1) // S.BLOCK)
        result = {"fitness": fit, "range": (lo, hi)}
    elif mode == "search":
        P, K = int(sys.argv[3]), int(sys.argv[4])
        pairs = W.protein_pairs(0, P, 40, 70)
        mats = W.candidate_matrices(K)
        freqs = W.pair_frequencies(pairs, 24)
        fit, (lo, hi), bk, bf, f, st = S.matrix_search_sharded(pairs, mats, 5, 5, -1.0, 576.0, freqs, device=0)
        result = {"fitness": fit, "range": (lo, hi), "best_k": bk, "f": f}
    elif mode == "batch":
        P = int(sys.argv[3])
        pairs = W.dna_read_batch(0, P, n=150)
        got = S.align_batch_sharded(ab.LOCAL, pairs, ab.dna_matrix(5, -4), 10, 2, device=0)
        cgot = S.align_batch_sharded(ab.LOCAL, pairs, ab.dna_matrix(5, -4), 10, 2, device=0, compact=True)
        if rank == 0:
            result = []
            for (lo, hi, r), (clo, chi, c) in zip(got, cgot):
                assert (lo, hi) == (clo, chi)
                ex = c.expand()
                strings = [r.alignment(i) for i in (0, r.count // 2, r.count - 1)]
                cstrings = [ex.alignment(i) for i in (0, r.count // 2, r.count - 1)]
                same = all((a[0] == b[0]).all() and (a[1] == b[1]).all() for a, b in zip(strings, cstrings))
                result.append((lo, hi, r.f, r.coords, r.aln_len, bool(same and (ex.coords == r.coords).all()), strings))
    elif mode == "gather":  # host logic only (CPU test): raw-tensor gather of ragged arrays
        rng = np.random.default_rng(100 + rank)
        arrs = [rng.normal(size=5 + 3 * rank), rng.integers(0, 200, (rank * 4, 4)).astype(np.uint64),
                rng.integers(0, 255, 1000 * (world - rank)).astype(np.uint8), np.zeros(0, dtype=np.uint32)]
        got = S._gather_arrays(arrs, dst=0)
        result = {"sent": arrs, "got": got}
    elif mode == "windows":  # host logic only (CPU test): the ranges tile the windows
        count = int(sys.argv[3])
        result = S.window_range(count, rank, world)
    elif mode == "pwm":
        length = int(sys.argv[3])
        seq, repeat, _ = W.dna_chromosome_with_repeats(length, 80, 12, seed=2)
        pwm = W.pwm_from_repeat(repeat, 1.0, -1.0) + 0.25
        result = S.pwm_scan_sharded(seq, pwm, 3.0, 1.0, first=5, step=20, window=100, device=0)
    elif mode == "scores":
        P = int(sys.argv[3])
        pairs = W.protein_pairs(0, P, 40, 90)
        result = S.score_batch_sharded(ab.LOCAL, pairs, ab.get_blosum62(), 11, 2, device=0)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(result, gathered, dst=0)
    if rank == 0:
        with open(out_path, "wb") as fh:
            pickle.dump(gathered, fh)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
