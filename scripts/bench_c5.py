#!/usr/bin/env python
"""BASELINE config 5 (heuristic matrix search) on N GPUs: pairs sharded across ranks, one NCCL
all-reduce of the blocked fitness sums.

  python scripts/bench_c5.py --pairs 2000 --mats 100                      (1 GPU)
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_c5.py ...

One JSON line from rank 0: GCUPS = sum over (pair, matrix) items of passes * n * m / wall seconds of
the sharded call (host buffers in, fitness out, max over ranks), FP64."""
import argparse
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import aligner_b200 as ab  # noqa: E402
from aligner_b200 import sharding as S  # noqa: E402
from aligner_b200 import workloads as W  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=2000)
    ap.add_argument("--mats", type=int, default=100)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pairs = W.protein_pairs(0, a.pairs, 270, 330)
    mats = W.candidate_matrices(a.mats)
    freqs = W.pair_frequencies(pairs, 24)
    dev = torch.device("cuda", local)
    S.matrix_search_sharded(S.slice_pairs(pairs, 0, min(a.pairs, 8)), mats[:2], 5, 5, -1.0, 576.0,
                            freqs[:min(a.pairs, 8)], device=local, collective_device=dev)  # warm-up
    best = None
    for _ in range(a.reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fit, (lo, hi), bk, bf, f, st = S.matrix_search_sharded(pairs, mats, 5, 5, -1.0, 576.0, freqs,
                                                               device=local, collective_device=dev)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        best = float(dt[0]) if best is None else min(best, float(dt[0]))
    # cells need the pass counts: a second (untimed) call on this rank's shard
    sub = S.slice_pairs(pairs, lo, hi)
    res = ab.heuristic_batch(sub, mats, 5, 5, -1.0, 576.0, freqs[lo:hi], want_matrices=False, device=local)
    n = np.diff(sub.q_off).astype(np.float64)
    m = np.diff(sub.t_off).astype(np.float64)
    cells = torch.tensor([float((res.passes.reshape(hi - lo, a.mats) * (n * m)[:, None]).sum())],
                         dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(cells)
    if rank == 0:
        print(json.dumps({
            "config": "C5 heuristic matrix search, %d pairs (270-330 aa) x %d candidate matrices, FP64" % (a.pairs, a.mats),
            "n_gpus": world, "seconds": best, "cells": float(cells[0]), "gcups": float(cells[0]) / best / 1e9,
            "fitness_sha256_16": hashlib.sha256(fit.tobytes()).hexdigest()[:16],
            "collective": "one NCCL all-reduce of [%d blocks x %d] f64 block sums" % ((a.pairs + S.BLOCK - 1) // S.BLOCK, a.mats),
        }), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
