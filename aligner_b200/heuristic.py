"""Heuristic matrix search over the C ABI.

Mirrors ``HeuristicAligner`` (aligner-core/src/heuristic/mod.rs:10-78),
``aligner_helpers::matrices::transform_matrix`` (aligner-helpers/src/matrices/mod.rs:19-68) and the
outer search the web layer spreads over Kafka workers (aligner-web/src/bin/dispatcher/handlers.rs:
193-251, aligner-web/src/bin/node/main.rs:61-74).
"""
import ctypes as C

import numpy as np

from . import _native as N
from .aligners import Heuristics, SimpleLocalAligner, _as_codes, _raise_status
from .batch import PackedPairs
from .enums import Error, Protein


def transform_matrix(matrix, k_d, r_squared, frequencies):
    """aligner_helpers::matrices::transform_matrix; raises Error('WrongMatrixSpecified')."""
    m = np.ascontiguousarray(matrix, dtype=np.float64)
    fr = np.ascontiguousarray(frequencies, dtype=np.float64)
    assert m.ndim == 2 and fr.shape == (m.shape[0],)
    out = np.empty_like(m)
    st = N.lib().alnr_transform_matrix(m.ctypes.data_as(N.f64p), m.shape[0], m.shape[1], float(k_d),
                                       float(r_squared), fr.ctypes.data_as(N.f64p), out.ctypes.data_as(N.f64p))
    if st == 9:
        raise Error("WrongMatrixSpecified")
    N.check(st)
    return out


class HeuristicBatchResult:
    def __init__(self, items, V, want_matrices):
        self.f = np.zeros(items, dtype=np.float64)
        self.passes = np.zeros(items, dtype=np.int32)
        self.status = np.zeros(items, dtype=np.int32)
        self.coords = np.zeros((items, 4), dtype=np.uint64)
        self.matrices = np.zeros((items, V, V), dtype=np.float64) if want_matrices else None


def heuristic_batch(pairs, matrices, del_, ext, kd, r_squared, frequencies, pair_idx=None, mat_idx=None,
                    want_matrices=True, device=-1):
    """HeuristicAligner::perform_alignment over many (pair, candidate matrix) items.

    pair_idx/mat_idx None = the full grid, item = pair * K + k."""
    mats = np.ascontiguousarray(matrices, dtype=np.float64)
    K, V = mats.shape[0], mats.shape[1]
    fr = np.ascontiguousarray(frequencies, dtype=np.float64)
    assert fr.shape == (pairs.count, V)
    if pair_idx is None:
        items = pairs.count * K
        pi = mi = None
    else:
        pi = np.ascontiguousarray(pair_idx, dtype=np.uint32)
        mi = np.ascontiguousarray(mat_idx, dtype=np.uint32)
        items = len(pi)
    res = HeuristicBatchResult(items, V, want_matrices)
    dummy = np.zeros((V, V))
    p = N.Params(N.LOCAL, V, float(del_), float(ext), dummy.ctypes.data_as(N.f64p), device, N.NUM_AUTO)
    it = N.HeuristicItems(pi.ctypes.data_as(N.u32p) if pi is not None else N.u32p(),
                          mi.ctypes.data_as(N.u32p) if mi is not None else N.u32p(), items,
                          mats.ctypes.data_as(N.f64p), K, float(kd), float(r_squared), fr.ctypes.data_as(N.f64p))
    out = N.HeuristicOut(res.f.ctypes.data_as(N.f64p), res.passes.ctypes.data_as(N.i32p),
                         res.status.ctypes.data_as(N.i32p), res.coords.ctypes.data_as(N.u64p),
                         res.matrices.ctypes.data_as(N.f64p) if want_matrices else N.f64p())
    b = pairs.c_struct()
    N.check(N.lib().alnr_heuristic_batch(C.byref(p), C.byref(b), C.byref(it), C.byref(out)))
    return res


def matrix_search(pairs, matrices, del_, ext, kd, r_squared, frequencies, device=-1, want_grid=True, n_gpus=None):
    """All pairs x all candidate matrices -> (fitness[K], best_k[P], best_f[P], f_grid[P,K], status[P,K]).
    ``n_gpus``: shard the pairs over that many devices inside the library (alnr_matrix_search_multi)."""
    mats = np.ascontiguousarray(matrices, dtype=np.float64)
    K, V = mats.shape[0], mats.shape[1]
    fr = np.ascontiguousarray(frequencies, dtype=np.float64)
    P = pairs.count
    fitness = np.zeros(K)
    best_k = np.zeros(P, dtype=np.uint32)
    best_f = np.zeros(P)
    grid = np.zeros((P, K)) if want_grid else None
    sgrid = np.zeros((P, K), dtype=np.int32) if want_grid else None
    dummy = np.zeros((V, V))
    p = N.Params(N.LOCAL, V, float(del_), float(ext), dummy.ctypes.data_as(N.f64p), device, N.NUM_AUTO)
    b = pairs.c_struct()
    args = (C.byref(p), C.byref(b), mats.ctypes.data_as(N.f64p), K, float(kd), float(r_squared),
            fr.ctypes.data_as(N.f64p), fitness.ctypes.data_as(N.f64p), best_k.ctypes.data_as(N.u32p),
            best_f.ctypes.data_as(N.f64p), grid.ctypes.data_as(N.f64p) if want_grid else N.f64p(),
            sgrid.ctypes.data_as(N.i32p) if want_grid else N.i32p())
    if n_gpus is None:
        N.check(N.lib().alnr_matrix_search(*args))
    else:
        N.check(N.lib().alnr_matrix_search_multi(*args, int(n_gpus)))
    return fitness, best_k, best_f, grid, sgrid


class HeuristicAligner:
    """aligner-core/src/heuristic/mod.rs:10-78."""

    def __init__(self, symbols, query, target):
        self.symbols = symbols
        self.query = query
        self.target = target

    @classmethod
    def from_str_seqs(cls, query, target, symbols=Protein):
        return cls(symbols, symbols.str_to_vec(query), symbols.str_to_vec(target))

    @classmethod
    def from_seqs(cls, query, target, symbols=Protein):
        return cls(symbols, _as_codes(query).copy(), _as_codes(target).copy())

    def perform_alignment(self, del_, ext, matrix, heuristics=None, want_matrices=True, device=-1):
        if heuristics is None:
            raise Error("MissingArgument")  # heuristic/mod.rs:43-46
        assert isinstance(heuristics, Heuristics)
        V = self.symbols.volume()
        mat = np.ascontiguousarray(matrix, dtype=np.float64)
        if mat.shape != (V, V):
            raise Error("MatrixShapeError")
        pairs = PackedPairs.from_lists([self.query], [self.target])
        res = heuristic_batch(pairs, mat[None], del_, ext, heuristics.kd, heuristics.r_squared,
                              heuristics.frequencies[None], want_matrices=True, device=device)
        if res.status[0] != 0:
            _raise_status(int(res.status[0]))
        # The loop returns its LAST pass (heuristic/mod.rs:73-76); re-running that pass with the
        # returned matrix reproduces it with the full AlignmentResult (strings, H, D).
        final = res.matrices[0]
        out = SimpleLocalAligner(self.symbols, self.query, self.target).perform_alignment(
            del_, ext, final, None, want_matrices=want_matrices, device=device)
        assert out.alignment.f == res.f[0]
        out.matrix = final
        out.passes = int(res.passes[0])
        return out
