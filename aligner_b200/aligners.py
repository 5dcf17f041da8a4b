"""Host-side mirror of the reference's alignment API over the C ABI.

Same names, argument meaning and error behaviour as aligner-core/src/lib.rs:21-59
(``Heuristics``, ``AlignerTrait``, ``AlignmentTrait``), aligner-core/src/simple/mod.rs
(``SimpleGlobalAligner`` :9-12, ``SimpleLocalAligner`` :14-17), aligner-core/src/alignment.rs
(``Alignment`` :4-10) and aligner-core/src/alignment_result.rs:6-13 (``AlignmentResult``).
All DP work runs in the CUDA library; nothing here computes a DP cell.
"""
import ctypes as C

import numpy as np

from . import _native as N
from .enums import Error, Protein


class Heuristics:
    """aligner-core/src/lib.rs:21-25"""

    def __init__(self, kd, r_squared, frequencies):
        self.kd = float(kd)
        self.r_squared = float(r_squared)
        self.frequencies = np.ascontiguousarray(frequencies, dtype=np.float64)


class Alignment:
    """aligner-core/src/alignment.rs:4-10 (+ AlignmentTrait impl :12-43)"""

    def __init__(self, symbols, query, target, coords, f):
        self.symbols = symbols
        self.query = query
        self.target = target
        self.coords = coords
        self.f = f

    def get_frequency_matrix(self):
        V = self.symbols.volume()
        F = np.zeros((V, V), dtype=np.float64)
        keep = (self.query != self.symbols.blank()) & (self.target != self.symbols.blank())
        np.add.at(F, (self.target[keep].astype(np.intp), self.query[keep].astype(np.intp)), 1.0)
        return F

    def get_alignment(self, matrix):
        q, t = self.query, self.target
        blank, pos = self.symbols.blank(), self.symbols.pos()
        out = np.full(len(q), blank, dtype=np.uint8)
        same = q == t
        out[same] = q[same]
        both = (~same) & (q != blank) & (t != blank)
        idx = np.nonzero(both)[0]
        ok = matrix[t[idx].astype(np.intp), q[idx].astype(np.intp)] >= 0.0
        out[idx[ok]] = pos
        return out

    def strings(self):
        return self.symbols.vec_to_str(self.query), self.symbols.vec_to_str(self.target)


class AlignmentResult:
    """aligner-core/src/alignment_result.rs:6-13"""

    def __init__(self, alignment_matrix, direction_matrix, alignment, matrix=None, numeric=0):
        self.alignment_matrix = alignment_matrix
        self.direction_matrix = direction_matrix
        self.alignment = alignment
        self.matrix = matrix
        self.numeric = numeric  # which device arithmetic ran (I16X2 / I32 / F64)


def _as_codes(seq):
    a = np.ascontiguousarray(seq, dtype=np.uint8)
    if a.ndim != 1:
        raise Error("ValidationError", "sequence must be one-dimensional")
    return a


def _raise_status(status):
    # aligner_core::Error variants keep their names; panics of the reference become Errors too
    name = N.STATUS_NAMES.get(status, str(status))
    if name in Error.KINDS:
        raise Error(name)
    raise N.AlignerError(status, (N.lib().alnr_last_error_detail() or b"").decode())


class _SimpleAligner:
    MODE = None

    def __init__(self, symbols, query, target):
        self.symbols = symbols
        self.query = query
        self.target = target

    @classmethod
    def from_str_seqs(cls, query, target, symbols=Protein):
        """AlignerTrait::from_str_seqs (simple/mod.rs:22-33, :148-159)"""
        return cls(symbols, symbols.str_to_vec(query), symbols.str_to_vec(target))

    @classmethod
    def from_seqs(cls, query, target, symbols=Protein):
        """AlignerTrait::from_seqs (simple/mod.rs:35-40, :161-166): copies the sequences."""
        return cls(symbols, _as_codes(query).copy(), _as_codes(target).copy())

    def perform_alignment(self, del_, ext, matrix, heuristics=None, want_matrices=True, device=-1,
                          numeric=N.NUM_AUTO):
        """AlignerTrait::perform_alignment (simple/mod.rs:42-144 global, :168-264 local)."""
        if heuristics is not None:
            raise Error("UnnecessaryArgument")  # simple/mod.rs:49-51, :175-177
        V = self.symbols.volume()
        mat = np.ascontiguousarray(matrix, dtype=np.float64)
        if mat.shape != (V, V):
            raise Error("MatrixShapeError", "expected %dx%d" % (V, V))
        q, t = self.query, self.target
        p = N.Params(self.MODE, V, float(del_), float(ext), mat.ctypes.data_as(N.f64p), device, numeric)
        res = N.PairResult()
        st = N.lib().alnr_align_pair(C.byref(p), q.ctypes.data_as(N.u8p), len(q),
                                     t.ctypes.data_as(N.u8p), len(t), 1 if want_matrices else 0,
                                     C.byref(res))
        if st != N.OK:
            _raise_status(st)
        try:
            L = int(res.aln_len)
            qa = np.ctypeslib.as_array(res.q_aln, shape=(L,)).copy()
            ta = np.ctypeslib.as_array(res.t_aln, shape=(L,)).copy()
            H = D = None
            if want_matrices:
                shape = (len(t) + 1, len(q) + 1)
                H = np.ctypeslib.as_array(res.H, shape=shape).copy()
                D = np.ctypeslib.as_array(res.D, shape=shape).copy()
            c = res.coords
            aln = Alignment(self.symbols, qa, ta, ((int(c[0]), int(c[1])), (int(c[2]), int(c[3]))),
                            float(res.f))
            out = AlignmentResult(H, D, aln, None, int(res.numeric))
            out.corner_score = float(res.corner_score)
            return out
        finally:
            N.lib().alnr_free_pair_result(C.byref(res))


class SimpleGlobalAligner(_SimpleAligner):
    """Needleman-Wunsch as the reference computes it (simple/mod.rs:42-144)."""
    MODE = N.GLOBAL


class SimpleLocalAligner(_SimpleAligner):
    """The reference's zero-floor-free local alignment (simple/mod.rs:168-264)."""
    MODE = N.LOCAL
