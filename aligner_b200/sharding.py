"""Multi-GPU layer: one process per GPU, independent shards, no data-path collective.

The reference spreads its work as one stateless job per (pair, matrix) item over Kafka workers
(aligner-web/src/bin/node/main.rs:61-74; dispatcher/handlers.rs:193-251).  Here every rank takes a
contiguous range of pair indices; batched alignment needs only a final gather of the results,
and the matrix search needs ONE reduction, of the fitness vector (``torch.distributed`` over NCCL
on GPUs, gloo in the CPU tests).  The reduction is made in a fixed blocked order
(``alnr_fitness_block_sums``) so the f64 sums are bit-identical for every number of GPUs.
"""
import ctypes as C

import numpy as np

from . import _native as N
from .batch import BatchResult, PackedPairs, align_batch
from .heuristic import heuristic_batch

BLOCK = N.FITNESS_BLOCK


def shard_bounds(weights, world, align=1):
    """Cut ``len(weights)`` consecutive items into ``world`` contiguous ranges of nearly equal total
    weight (balanced by sum of m*n); every boundary is a multiple of ``align``.
    Returns world+1 boundaries."""
    w = np.asarray(weights, dtype=np.float64)
    count = len(w)
    cum = np.concatenate([[0.0], np.cumsum(w)])
    bounds = [0]
    for r in range(1, world):
        target = cum[-1] * r / world
        i = int(np.searchsorted(cum, target, side="left"))
        i = int(round(i / align)) * align
        bounds.append(max(min(i, count // align * align), bounds[-1]))
    bounds.append(count)
    return bounds


def shard_range(count, rank, world, weights=None, align=1):
    """[lo, hi) of rank ``rank``."""
    b = shard_bounds(np.ones(count) if weights is None else weights, world, align)
    return b[rank], b[rank + 1]


def slice_pairs(pairs, lo, hi):
    """Pairs [lo, hi) of a PackedPairs as a new PackedPairs (views, offsets rebased)."""
    q0, q1 = int(pairs.q_off[lo]), int(pairs.q_off[hi])
    t0, t1 = int(pairs.t_off[lo]), int(pairs.t_off[hi])
    return PackedPairs(pairs.q[q0:q1], pairs.q_off[lo:hi + 1] - np.uint64(q0),
                       pairs.t[t0:t1], pairs.t_off[lo:hi + 1] - np.uint64(t0))


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _rank_world():
    d = _dist()
    return (d.get_rank(), d.get_world_size()) if d else (0, 1)


def fitness_block_sums(f_grid, status_grid=None):
    """alnr_fitness_block_sums: [ceil(P/BLOCK), K] partial sums of a local f grid (pure host math)."""
    f = np.ascontiguousarray(f_grid, dtype=np.float64)
    P, K = f.shape
    st = None if status_grid is None else np.ascontiguousarray(status_grid, dtype=np.int32)
    out = np.zeros(((P + BLOCK - 1) // BLOCK, K), dtype=np.float64)
    N.check(N.lib().alnr_fitness_block_sums(f.ctypes.data_as(N.f64p),
                                            st.ctypes.data_as(N.i32p) if st is not None else N.i32p(),
                                            P, K, out.ctypes.data_as(N.f64p)))
    return out


def fitness_from_blocks(block_sums):
    b = np.ascontiguousarray(block_sums, dtype=np.float64)
    out = np.zeros(b.shape[1], dtype=np.float64)
    N.check(N.lib().alnr_fitness_from_blocks(b.ctypes.data_as(N.f64p), b.shape[0], b.shape[1],
                                             out.ctypes.data_as(N.f64p)))
    return out


def reduce_fitness(local_blocks, first_block, total_blocks, device=None):
    """The one collective of the search: every rank contributes the blocks it owns to a zero
    [total_blocks, K] tensor, one all-reduce(SUM) (each block has a single owner, so adding the
    other ranks' zeros is exact), then every rank adds the blocks in block order."""
    import torch
    K = local_blocks.shape[1]
    full = np.zeros((total_blocks, K), dtype=np.float64)
    full[first_block:first_block + local_blocks.shape[0]] = local_blocks
    d = _dist()
    if d is not None and d.get_world_size() > 1:
        t = torch.from_numpy(full)
        if device is not None:
            t = t.to(device)
        d.all_reduce(t, op=d.ReduceOp.SUM)
        full = t.cpu().numpy()
    return fitness_from_blocks(full)


def matrix_search_sharded(pairs, matrices, del_, ext, kd, r_squared, frequencies, device=-1,
                          collective_device=None):
    """C5 on N ranks: shard the PAIRS on fitness-block boundaries (balanced by sum of n*m), every
    rank runs the heuristic loop for its pairs under all K candidate matrices on its own GPU, the
    fitness vector is reduced across ranks.  Returns (fitness[K] identical on every rank,
    (lo, hi), best_k[lo:hi], best_f[lo:hi], f_grid[lo:hi], status[lo:hi])."""
    rank, world = _rank_world()
    mats = np.ascontiguousarray(matrices, dtype=np.float64)
    K = mats.shape[0]
    P = pairs.count
    cells = np.diff(pairs.q_off).astype(np.float64) * np.diff(pairs.t_off).astype(np.float64)
    lo, hi = shard_range(P, rank, world, cells, align=BLOCK)
    total_blocks = (P + BLOCK - 1) // BLOCK
    if hi > lo:
        res = heuristic_batch(slice_pairs(pairs, lo, hi), mats, del_, ext, kd, r_squared,
                              np.ascontiguousarray(frequencies[lo:hi]), want_matrices=False, device=device)
        f = res.f.reshape(hi - lo, K)
        st = res.status.reshape(hi - lo, K)
        blocks = fitness_block_sums(f, st)
    else:
        f = np.zeros((0, K))
        st = np.zeros((0, K), dtype=np.int32)
        blocks = np.zeros((0, K))
    fitness = reduce_fitness(blocks, lo // BLOCK, total_blocks, collective_device)
    masked = np.where(st == 0, f, -np.inf)
    best_k = masked.argmax(axis=1).astype(np.uint32) if hi > lo else np.zeros(0, dtype=np.uint32)
    best_f = masked.max(axis=1) if hi > lo else np.zeros(0)
    return fitness, (lo, hi), best_k, best_f, f, st


def _gather_arrays(arrs, dst=0):
    """Rank ``dst`` receives every rank's list of numpy arrays, the others get None.  No pickling of
    the payload: the byte sizes travel in one all_gather of an int64 tensor, the arrays as ONE uint8
    tensor per rank (padded to the longest) through ``gather`` — on the GPU for NCCL, on the CPU for
    gloo."""
    import torch
    d = _dist()
    rank, world = _rank_world()
    dev = torch.device("cuda", torch.cuda.current_device()) if d.get_backend() == "nccl" else torch.device("cpu")
    arrs = [np.ascontiguousarray(a) for a in arrs]
    sizes = torch.tensor([a.nbytes for a in arrs], dtype=torch.int64, device=dev)
    all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
    d.all_gather(all_sizes, sizes)
    totals = [int(t.sum()) for t in all_sizes]
    cap = max(max(totals), 1)
    buf = torch.zeros(cap, dtype=torch.uint8, device=dev)
    pos = 0
    for a in arrs:
        if a.nbytes:
            buf[pos:pos + a.nbytes] = torch.from_numpy(a.reshape(-1).view(np.uint8)).to(dev)
        pos += a.nbytes
    bins = [torch.zeros(cap, dtype=torch.uint8, device=dev) for _ in range(world)] if rank == dst else None
    d.gather(buf, bins, dst=dst)
    if rank != dst:
        return None
    out = []
    for r in range(world):
        host = bins[r].cpu().numpy()
        pos, got = 0, []
        for k, a in enumerate(arrs):  # same dtypes and trailing shapes on every rank
            nb = int(all_sizes[r][k])
            got.append(host[pos:pos + nb].view(a.dtype).reshape((-1,) + a.shape[1:]).copy())
            pos += nb
        out.append(got)
    return out


def align_batch_sharded(mode, pairs, matrix, del_, ext, device=-1, gather=True, compact=False):
    """C3 on N ranks: this rank aligns its contiguous range of pairs (balanced by sum of n*m) on its
    own GPU; no collective on the data path.  With ``gather`` rank 0 receives every rank's
    (lo, hi, result) in rank order, the other ranks get their own only; the arrays travel as raw
    tensors (``_gather_arrays``), nothing is pickled.  ``compact``: results in the compact format
    (CompactResult: start cell + 2-bit moves, ~5x fewer bytes to gather).  Inside one process
    ``align_batch(..., n_gpus=N)`` (alnr_align_batch_multi) does the same without any gather."""
    from .batch import CompactResult
    rank, world = _rank_world()
    cells = np.diff(pairs.q_off).astype(np.float64) * np.diff(pairs.t_off).astype(np.float64)
    lo, hi = shard_range(pairs.count, rank, world, cells)
    sub = slice_pairs(pairs, lo, hi)
    res = align_batch(mode, sub, matrix, del_, ext, device=device, compact=compact)
    mine = (lo, hi, res)
    d = _dist()
    if not gather or d is None or world == 1:
        return [mine]
    if compact:
        used = res.move_words()
        fields = ["f", "corner", "coords", "end_yx", "aln_len", "status", "moves_off", "moves"]
        arrs = [getattr(res, k) if k != "moves" else res.moves[:used] for k in fields]
    else:
        fields = ["f", "corner", "coords", "aln_off", "aln_len", "status", "q_aln", "t_aln"]
        arrs = [getattr(res, k) for k in fields]
    bounds = np.array([lo, hi], dtype=np.int64)
    got = _gather_arrays([bounds] + arrs, dst=0)
    if rank != 0:
        return [mine]
    out = []
    for r in range(world):
        rlo, rhi = (int(v) for v in got[r][0])
        if compact:
            rr = CompactResult.__new__(CompactResult)
            rr.pairs, rr.mode, rr.volume = slice_pairs(pairs, rlo, rhi), mode, np.asarray(matrix).shape[0]
        else:
            rr = BatchResult.__new__(BatchResult)
            rr.end_yx = None
        rr.count = rhi - rlo
        for k, arr in zip(fields, got[r][1:]):
            setattr(rr, k, arr)
        out.append((rlo, rhi, rr))
    return out


def score_batch_sharded(mode, pairs, matrix, del_, ext, device=-1, gather=True):
    """N2 on N ranks: every rank scores its contiguous range of pairs (balanced by sum of n*m); the
    8-byte scores are all-gathered so that every rank can run the Gumbel fit.  Returns (f[count],
    status[count]) on every rank when ``gather`` (else the rank's own slice and (lo, hi))."""
    from .batch import score_batch
    rank, world = _rank_world()
    cells = np.diff(pairs.q_off).astype(np.float64) * np.diff(pairs.t_off).astype(np.float64)
    lo, hi = shard_range(pairs.count, rank, world, cells)
    res = score_batch(mode, slice_pairs(pairs, lo, hi), matrix, del_, ext, device=device)
    d = _dist()
    if not gather or d is None or world == 1:
        return (res.f, res.status) if gather else (res.f, res.status, (lo, hi))
    parts = [None] * world
    d.all_gather_object(parts, (lo, hi, res.f, res.status))
    f = np.zeros(pairs.count)
    st = np.zeros(pairs.count, dtype=np.int32)
    for plo, phi, pf, pst in parts:
        f[plo:phi] = pf
        st[plo:phi] = pst
    return f, st


def window_range(count, rank, world):
    """Windows [lo, hi) of rank ``rank``: equal counts (the windows of a scan have equal sizes)."""
    return count * rank // world, count * (rank + 1) // world


def pwm_scan_sharded(seq, pwm, del_, ext, first=0, step=1, window=0, device=-1, resident=None):
    """N1 on N ranks: every rank keeps the whole sequence resident in its own HBM (1 byte per base)
    and scans a contiguous range of the strided windows score-only; the f values are all-gathered (8
    bytes per window — no collective on the data path).  Returns f[count] on every rank."""
    from .pwm import ResidentSequence
    rank, world = _rank_world()
    n = len(seq) if resident is None else resident.len
    count = 0 if first >= n else (n - first + step - 1) // step
    lo, hi = window_range(count, rank, world)
    rs = resident if resident is not None else ResidentSequence(seq, device)
    try:
        mine = rs.align_windows(pwm, del_, ext, first=first + lo * step, step=step, window=window, count=hi - lo).f
    finally:
        if resident is None:
            rs.close()
    d = _dist()
    if d is None or world == 1:
        return mine
    parts = [None] * world
    d.all_gather_object(parts, (lo, hi, mine))
    f = np.zeros(count)
    for plo, phi, pf in parts:
        f[plo:phi] = pf
    return f
