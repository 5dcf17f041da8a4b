"""Batched alignment of independent pairs over the C ABI (alnr_align_batch / alnr_batch_*).

The reference aligns batches by looping ``perform_alignment`` on OS threads
(aligner-core/src/statistics/mod.rs:255-286; latent-repeat-search/engine/calc.rs:47-69);
here one call hands the whole batch to the GPU.
"""
import ctypes as C

import numpy as np

from . import _native as N


class PackedPairs:
    """Concatenated symbol codes + uint64 offsets for queries and targets."""

    def __init__(self, q, q_off, t, t_off):
        self.q = np.ascontiguousarray(q, dtype=np.uint8)
        self.t = np.ascontiguousarray(t, dtype=np.uint8)
        self.q_off = np.ascontiguousarray(q_off, dtype=np.uint64)
        self.t_off = np.ascontiguousarray(t_off, dtype=np.uint64)
        self.count = len(self.q_off) - 1
        assert len(self.t_off) == self.count + 1

    @classmethod
    def from_lists(cls, queries, targets):
        assert len(queries) == len(targets)
        qo = np.zeros(len(queries) + 1, dtype=np.uint64)
        to = np.zeros(len(targets) + 1, dtype=np.uint64)
        qo[1:] = np.cumsum([len(x) for x in queries])
        to[1:] = np.cumsum([len(x) for x in targets])
        cat = lambda xs: (np.concatenate([np.asarray(x, dtype=np.uint8) for x in xs])
                          if len(xs) else np.zeros(0, dtype=np.uint8))
        return cls(cat(queries), qo, cat(targets), to)

    @classmethod
    def uniform(cls, q2d, t2d):
        """count x n and count x m code arrays (all pairs the same shape)."""
        q2d = np.ascontiguousarray(q2d, dtype=np.uint8)
        t2d = np.ascontiguousarray(t2d, dtype=np.uint8)
        c, n = q2d.shape
        m = t2d.shape[1]
        return cls(q2d.reshape(-1), np.arange(c + 1, dtype=np.uint64) * np.uint64(n),
                   t2d.reshape(-1), np.arange(c + 1, dtype=np.uint64) * np.uint64(m))

    def pair(self, i):
        return (self.q[int(self.q_off[i]):int(self.q_off[i + 1])],
                self.t[int(self.t_off[i]):int(self.t_off[i + 1])])

    def cells(self):
        n = np.diff(self.q_off).astype(np.float64)
        m = np.diff(self.t_off).astype(np.float64)
        return float((n * m).sum())

    def aln_capacity(self):
        return int(self.q_off[-1]) + int(self.t_off[-1]) + self.count

    def c_struct(self):
        return N.Batch(self.q.ctypes.data_as(N.u8p), self.q_off.ctypes.data_as(N.u64p),
                       self.t.ctypes.data_as(N.u8p), self.t_off.ctypes.data_as(N.u64p), self.count,
                       N.u32p(), N.u32p())

    def pack2bit(self):
        """The same batch in the compact input format (alnr_batch.q2 / .t2; DNA codes 0..3)."""
        return Packed2Pairs(pack2bit(self.q, self.q_off), self.q_off, pack2bit(self.t, self.t_off), self.t_off)


def pack2bit(codes, off):
    """2-bit packing of concatenated DNA codes: 16 codes per uint32, code k of a sequence at bits
    [2*(k%16), +2) of its word k//16, every sequence on a word boundary, sequences densely one
    after the other (include/aligner_b200.h, alnr_batch.q2)."""
    codes = np.asarray(codes, dtype=np.uint8)
    off = np.asarray(off, dtype=np.uint64).astype(np.int64)
    lens = np.diff(off)
    nw = (lens + 15) // 16
    woff = np.zeros(len(off), dtype=np.int64)
    woff[1:] = np.cumsum(nw)
    padded = np.zeros(int(woff[-1]) * 16, dtype=np.uint32)
    if len(codes):
        seq = np.repeat(np.arange(len(lens)), lens)            # sequence of every symbol
        k = np.arange(len(codes), dtype=np.int64) - off[seq]    # its index inside the sequence
        padded[woff[seq] * 16 + k] = codes & 3
    sh = (2 * np.arange(16, dtype=np.uint32))[None, :]
    return np.ascontiguousarray((padded.reshape(-1, 16) << sh).sum(axis=1, dtype=np.uint32))


class Packed2Pairs:
    """A batch in the compact input format: 2-bit DNA words + the symbol offsets (lengths)."""

    def __init__(self, q2, q_off, t2, t_off):
        self.q2 = np.ascontiguousarray(q2, dtype=np.uint32)
        self.t2 = np.ascontiguousarray(t2, dtype=np.uint32)
        self.q_off = np.ascontiguousarray(q_off, dtype=np.uint64)
        self.t_off = np.ascontiguousarray(t_off, dtype=np.uint64)
        self.count = len(self.q_off) - 1

    cells = PackedPairs.cells
    aln_capacity = PackedPairs.aln_capacity

    def c_struct(self):
        return N.Batch(N.u8p(), self.q_off.ctypes.data_as(N.u64p), N.u8p(), self.t_off.ctypes.data_as(N.u64p),
                       self.count, self.q2.ctypes.data_as(N.u32p), self.t2.ctypes.data_as(N.u32p))

    def bytes_in(self):
        return int(self.q2.nbytes + self.t2.nbytes)


class BatchResult:
    """Host-side outputs of one batched call (numpy arrays; alignments are slices)."""

    def __init__(self, pairs, want_alignments=True):
        c = pairs.count
        self.count = c
        self.f = np.zeros(c, dtype=np.float64)
        self.corner = np.zeros(c, dtype=np.float64)
        self.coords = np.zeros((c, 4), dtype=np.uint64)
        self.aln_off = np.zeros(c, dtype=np.uint64)
        self.aln_len = np.zeros(c, dtype=np.uint32)
        self.status = np.zeros(c, dtype=np.int32)
        cap = pairs.aln_capacity() if want_alignments else 0
        self.q_aln = np.zeros(cap, dtype=np.uint8) if want_alignments else None
        self.t_aln = np.zeros(cap, dtype=np.uint8) if want_alignments else None
        self.end_yx = None

    def c_struct(self):
        return N.BatchOut(_ptr(self.f, N.f64p), _ptr(self.corner, N.f64p), _ptr(self.coords, N.u64p),
                          _ptr(self.aln_off, N.u64p), _ptr(self.aln_len, N.u32p),
                          _ptr(self.q_aln, N.u8p), _ptr(self.t_aln, N.u8p), _ptr(self.status, N.i32p),
                          _ptr(getattr(self, "end_yx", None), N.u32p), N.u64p(), N.u32p())

    def alignment(self, i):
        o, l = int(self.aln_off[i]), int(self.aln_len[i])
        return self.q_aln[o:o + l], self.t_aln[o:o + l]

    def bytes_out(self):
        arrs = [self.f, self.corner, self.coords, self.aln_off, self.aln_len, self.status,
                self.q_aln, self.t_aln]
        return int(sum(a.nbytes for a in arrs if a is not None))


def _ptr(a, ty):
    return a.ctypes.data_as(ty) if a is not None else ty()


class CompactResult:
    """Outputs of a batched call in the compact result format (alnr_batch_out.moves): per pair the
    traceback's start cell, its length and one 2-bit Direction per step instead of two byte strings.
    ``Alignment.query`` / ``.target`` are private to aligner-core (alignment.rs:6-7); what callers
    can see — coords, f, the frequency matrix, the consensus line — is derived on demand by
    ``alignment`` / ``expand`` (alnr_expand_alignments on the host cores)."""

    def __init__(self, pairs, mode, volume=4, want_corner=False, want_coords=False, alloc=None):
        alloc = alloc or (lambda shape, dt: np.zeros(shape, dtype=dt))
        c = pairs.count
        self.pairs, self.mode, self.volume, self.count = pairs, mode, volume, c
        self.f = alloc((c,), np.float64)
        self.corner = alloc((c,), np.float64) if want_corner else None
        self.coords = alloc((c, 4), np.uint64) if want_coords else None
        self.end_yx = alloc((c, 2), np.uint32)
        self.aln_len = alloc((c,), np.uint32)
        self.status = alloc((c,), np.int32)
        self.moves_off = alloc((c,), np.uint64)
        b = pairs.c_struct()
        self.moves = alloc((int(N.lib().alnr_moves_capacity(C.byref(b))),), np.uint32)

    def c_struct(self):
        return N.BatchOut(_ptr(self.f, N.f64p), _ptr(self.corner, N.f64p), _ptr(self.coords, N.u64p), N.u64p(),
                          _ptr(self.aln_len, N.u32p), N.u8p(), N.u8p(), _ptr(self.status, N.i32p),
                          _ptr(self.end_yx, N.u32p), _ptr(self.moves_off, N.u64p), _ptr(self.moves, N.u32p))

    def move_words(self):
        """uint32 words of move strings actually used (what crossed PCIe)."""
        return int(((self.aln_len.astype(np.int64) + 14) >> 4).sum())

    def bytes_out(self):
        arrs = [self.f, self.corner, self.coords, self.end_yx, self.aln_len, self.status, self.moves_off]
        return int(sum(a.nbytes for a in arrs if a is not None)) + 4 * self.move_words()

    def expand(self, first=0, count=None, want_freq=False, threads=0):
        """BatchResult of the pairs [first, first+count): strings, coords, aln_off / aln_len exactly
        as the string format returns them (alignments back to back from offset 0)."""
        count = self.count - first if count is None else count
        res = BatchResult.__new__(BatchResult)
        res.count = count
        res.f = self.f[first:first + count]
        res.corner = self.corner[first:first + count] if self.corner is not None else None
        res.status = self.status[first:first + count]
        res.end_yx = self.end_yx[first:first + count]
        res.coords = np.zeros((count, 4), dtype=np.uint64)
        res.aln_off = np.zeros(count, dtype=np.uint64)
        res.aln_len = np.zeros(count, dtype=np.uint32)
        total = int(self.aln_len[first:first + count].astype(np.int64).sum())
        res.q_aln = np.zeros(max(total, 1), dtype=np.uint8)
        res.t_aln = np.zeros(max(total, 1), dtype=np.uint8)
        res.freq = np.zeros((count, self.volume, self.volume)) if want_freq else None
        b, c, e = self.pairs.c_struct(), self.c_struct(), res.c_struct()
        N.check(N.lib().alnr_expand_alignments(self.mode, self.volume, C.byref(b), C.byref(c), first, count,
                                               C.byref(e), _ptr(res.freq, N.f64p), threads))
        return res

    def alignment(self, i):
        """(query, target) aligned code arrays of pair i, expanded on demand."""
        return self.expand(i, 1).alignment(0)


def make_params(mode, matrix, del_, ext, device=-1, numeric=N.NUM_AUTO):
    mat = np.ascontiguousarray(matrix, dtype=np.float64)
    assert mat.ndim == 2 and mat.shape[0] == mat.shape[1]
    p = N.Params(mode, mat.shape[0], float(del_), float(ext), mat.ctypes.data_as(N.f64p), device, numeric)
    p._keep = mat  # keep the matrix alive as long as the struct
    return p


def align_batch(mode, pairs, matrix, del_, ext, device=-1, numeric=N.NUM_AUTO, want_alignments=True,
                out=None, compact=False, n_gpus=None):
    """Host buffers in -> host buffers out (alnr_align_batch).  ``pairs``: PackedPairs (byte codes)
    or Packed2Pairs (2-bit DNA).  ``compact``: return a CompactResult (move strings).  ``n_gpus``:
    shard the batch over that many devices inside the library (alnr_align_batch_multi; 0 = all)."""
    p = make_params(mode, matrix, del_, ext, device, numeric)
    if out is not None:
        res = out
    elif compact:
        res = CompactResult(pairs, mode, np.asarray(matrix).shape[0], want_corner=True, want_coords=True)
    else:
        res = BatchResult(pairs, want_alignments)
    b = pairs.c_struct()
    o = res.c_struct()
    if n_gpus is None:
        N.check(N.lib().alnr_align_batch(C.byref(p), C.byref(b), C.byref(o)))
    else:
        N.check(N.lib().alnr_align_batch_multi(C.byref(p), C.byref(b), C.byref(o), int(n_gpus)))
    return res


class ScoreResult:
    """Outputs of a score-only call: what the p-value shuffles keep of every alignment
    (``alignment.f``, aligner-core/src/statistics/mod.rs:268-276) plus the corner score and the
    1-based (target, query) cell the traceback would have started from."""

    def __init__(self, count):
        self.count = count
        self.f = np.zeros(count, dtype=np.float64)
        self.corner = np.zeros(count, dtype=np.float64)
        self.end_yx = np.zeros((count, 2), dtype=np.uint32)
        self.status = np.zeros(count, dtype=np.int32)

    def c_args(self):
        return (self.f.ctypes.data_as(N.f64p), self.corner.ctypes.data_as(N.f64p),
                self.end_yx.ctypes.data_as(N.u32p), self.status.ctypes.data_as(N.i32p))

    def bytes_out(self):
        return int(self.f.nbytes + self.corner.nbytes + self.end_yx.nbytes + self.status.nbytes)


def score_batch(mode, pairs, matrix, del_, ext, device=-1, numeric=N.NUM_AUTO, out=None):
    """Scores without alignments (alnr_score_batch): same fill kernels, no direction scratch and no
    traceback."""
    p = make_params(mode, matrix, del_, ext, device, numeric)
    res = out if out is not None else ScoreResult(pairs.count)
    b = pairs.c_struct()
    N.check(N.lib().alnr_score_batch(C.byref(p), C.byref(b), *res.c_args()))
    return res


class DeviceBatch:
    """Device-resident batch: upload once, run many times (alnr_batch_*).  ``score_only`` builds
    the handle with alnr_batch_create_scores; read it with ``download_scores``."""

    def __init__(self, mode, pairs, matrix, del_, ext, device=-1, numeric=N.NUM_AUTO, score_only=False,
                 compact=False):
        self.pairs = pairs
        self.score_only = score_only
        self.compact = compact
        self.mode, self.volume = mode, np.asarray(matrix).shape[0]
        self._p = make_params(mode, matrix, del_, ext, device, numeric)
        self._h = C.c_void_p()
        b = pairs.c_struct()
        create = (N.lib().alnr_batch_create_scores if score_only else
                  N.lib().alnr_batch_create_compact if compact else N.lib().alnr_batch_create)
        N.check(create(C.byref(self._p), C.byref(b), C.byref(self._h)))

    def download_scores(self, out=None):
        res = out if out is not None else ScoreResult(self.pairs.count)
        N.check(N.lib().alnr_batch_download_scores(self._h, *res.c_args()))
        return res

    def run(self):
        N.check(N.lib().alnr_batch_run(self._h))

    def sync(self):
        N.check(N.lib().alnr_batch_sync(self._h))

    def run_timed(self, steps=1, profile_fill=True):
        """`steps` passes between CUDA events on the batch's own stream -> (total ms, fill ms)."""
        ms, fill = C.c_float(), C.c_float()
        N.check(N.lib().alnr_batch_run_timed(self._h, steps, C.byref(ms),
                                             C.byref(fill) if profile_fill else None))
        return ms.value, fill.value

    def download(self, want_alignments=True, out=None):
        if out is not None:
            res = out
        elif self.compact:
            res = CompactResult(self.pairs, self.mode, self.volume, want_corner=True, want_coords=True)
        else:
            res = BatchResult(self.pairs, want_alignments)
        o = res.c_struct()
        N.check(N.lib().alnr_batch_download(self._h, C.byref(o)))
        return res

    @property
    def numeric(self):
        return N.lib().alnr_batch_numeric(self._h)

    @property
    def rerun_count(self):
        return N.lib().alnr_batch_rerun_count(self._h)

    @property
    def stream(self):
        return N.lib().alnr_batch_stream(self._h)

    def close(self):
        if self._h:
            N.lib().alnr_batch_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
