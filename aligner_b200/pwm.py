"""PWM aligner over the C ABI.

Mirrors ``PWMAligner`` (aligner-core/src/pwm/mod.rs:9-126), ``PWMAlignment``
(aligner-core/src/alignment.rs:45-92), ``HeuristicPWMAligner`` (aligner-core/src/heuristic/mod.rs:
81-141) and ``get_random_pwm`` (aligner-core/src/lib.rs:92-96).  The reference's batch use of
the PWM aligner — sliding windows of one chromosome on OS threads
(bin/latent-repeat-search/engine/calc.rs:47-69, :107-137) — is ``ResidentSequence.align_windows``:
the sequence lives in HBM, one call aligns any set of its windows to one PWM.
"""
import ctypes as C

import numpy as np

from . import _native as N
from .aligners import AlignmentResult, _as_codes, _raise_status
from .enums import DNA, Error
from .heuristic import transform_matrix


def get_random_pwm(length, rng=None):
    """lib.rs:92-96: a 4 x length matrix of uniform draws from {-1, 0, 1}."""
    rng = np.random.default_rng() if rng is None else rng
    return rng.integers(-1, 2, (4, length)).astype(np.float64)


class PWMAlignment:
    """aligner-core/src/alignment.rs:45-52 (+ AlignmentTrait impl :54-79, ``empty`` :82-92)"""

    def __init__(self, symbols, numbered, query, dim, coords, f):
        self.symbols = symbols
        self.numbered = numbered
        self.query = query
        self.dim = dim
        self.coords = coords
        self.f = f

    @classmethod
    def empty(cls, symbols=DNA):
        return cls(symbols, np.zeros(0, dtype=np.uint32), np.zeros(0, dtype=np.uint8), 0, ((0, 0), (0, 0)), 0.0)

    def get_frequency_matrix(self):
        F = np.zeros((self.symbols.volume(), self.dim), dtype=np.float64)
        keep = (self.numbered != 0) & (self.query != self.symbols.blank())
        np.add.at(F, (self.query[keep].astype(np.intp), self.numbered[keep].astype(np.intp) - 1), 1.0)
        return F

    def get_alignment(self, _matrix=None):
        return np.where(self.numbered != 0, self.query, np.uint8(self.symbols.blank())).astype(np.uint8)


class PwmWindowsResult:
    """Host arrays of one ``align_windows`` call (alnr_pwm_out)."""

    def __init__(self, count, L, slot, want_alignments, want_freq):
        self.count, self.L, self.slot = count, L, slot
        self.f = np.zeros(count, dtype=np.float64)
        self.coords = np.zeros((count, 4), dtype=np.uint64)
        self.status = np.zeros(count, dtype=np.int32)
        self.aln_off = np.zeros(count, dtype=np.uint64)
        self.aln_len = np.zeros(count, dtype=np.uint32)
        self.numbered = np.zeros(count * slot, dtype=np.uint32) if want_alignments else None
        self.query = np.zeros(count * slot, dtype=np.uint8) if want_alignments else None
        self.freq = np.zeros((count, 4, L), dtype=np.float64) if want_freq else None
        self.numeric = 0

    def alignment(self, i, symbols=DNA):
        """PWMAlignment of window i."""
        o, l = int(self.aln_off[i]), int(self.aln_len[i])
        c = self.coords[i]
        return PWMAlignment(symbols, self.numbered[o:o + l].copy(), self.query[o:o + l].copy(), self.L,
                            ((int(c[0]), int(c[1])), (int(c[2]), int(c[3]))), float(self.f[i]))


class ResidentSequence:
    """A DNA sequence (codes 0..3) uploaded once (alnr_pwm_seq)."""

    def __init__(self, seq, device=-1):
        self._h = C.c_void_p()
        a = _as_codes(seq)
        self.len = len(a)
        N.check(N.lib().alnr_pwm_seq_create(a.ctypes.data_as(N.u8p), len(a), device, C.byref(self._h)))

    def align_windows(self, pwm, del_, ext, starts=None, lens=None, first=0, step=1, window=0, count=None,
                      want_alignments=False, want_freq=False, want_matrices=False, numeric=N.NUM_AUTO):
        """PWMAligner::perform_alignment for every window.  Explicit windows: ``starts`` / ``lens``.
        Strided: window i = seq[first + i*step, +window) clipped at the end of the sequence (the
        j..border windows of calc.rs:50-56).  Score-only unless alignments / freq are asked for."""
        mat = np.ascontiguousarray(pwm, dtype=np.float64)
        if mat.ndim != 2:
            raise Error("MatrixShapeError")
        rows, L = mat.shape
        if starts is not None:
            st_a = np.ascontiguousarray(starts, dtype=np.uint64)
            ln_a = np.ascontiguousarray(lens, dtype=np.uint32)
            assert st_a.shape == ln_a.shape
            count = len(st_a)
            m_max = int(ln_a.max()) if count else 0
            w = N.Windows(count, st_a.ctypes.data_as(N.u64p), ln_a.ctypes.data_as(N.u32p), 0, 0, 0)
        else:
            if count is None:
                count = 0 if first >= self.len else (self.len - first + step - 1) // step
            m_max = min(window, max(self.len - first, 0))
            w = N.Windows(count, N.u64p(), N.u32p(), first, step, window)
        slot = L + m_max
        res = PwmWindowsResult(count, L, slot, want_alignments, want_freq)
        H = D = None
        if want_matrices:
            H = np.zeros((m_max + 1, L + 1), dtype=np.float64)
            D = np.zeros((m_max + 1, L + 1), dtype=np.uint8)
        ptr = lambda a, ty: a.ctypes.data_as(ty) if a is not None else ty()
        p = N.PwmParams(float(del_), float(ext), mat.ctypes.data_as(N.f64p), rows, L, -1, numeric)
        o = N.PwmOut(ptr(res.f, N.f64p), ptr(res.coords, N.u64p), ptr(res.status, N.i32p),
                     ptr(res.aln_off, N.u64p), ptr(res.aln_len, N.u32p), ptr(res.numbered, N.u32p),
                     ptr(res.query, N.u8p), slot, ptr(res.freq, N.f64p), ptr(H, N.f64p), ptr(D, N.u8p))
        st = N.lib().alnr_pwm_align_windows(C.byref(p), self._h, C.byref(w), C.byref(o))
        if st != N.OK:
            _raise_status(st)
        res.numeric = N.lib().alnr_pwm_last_numeric()
        res.H, res.D = H, D
        return res

    def close(self):
        if self._h:
            N.lib().alnr_pwm_seq_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PWMAligner:
    """aligner-core/src/pwm/mod.rs:9-126: one query against one 4 x L matrix."""

    def __init__(self, symbols, query):
        self.symbols = symbols
        self.query = query

    @classmethod
    def from_str_seqs(cls, query, _target="", symbols=DNA):
        """pwm/mod.rs:14-21: the target argument is ignored."""
        return cls(symbols, symbols.str_to_vec(query))

    @classmethod
    def from_seqs(cls, query, _target=(), symbols=DNA):
        """pwm/mod.rs:23-27"""
        return cls(symbols, _as_codes(query).copy())

    def perform_alignment(self, del_, ext, matrix, heuristics=None, want_matrices=True, device=-1,
                          numeric=N.NUM_AUTO):
        """pwm/mod.rs:29-126"""
        if heuristics is not None:
            raise Error("UnnecessaryArgument")  # pwm/mod.rs:36-38
        mat = np.ascontiguousarray(matrix, dtype=np.float64)
        if mat.ndim != 2 or mat.shape[0] != 4:
            raise Error("MatrixShapeError")     # pwm/mod.rs:40-42
        seq = ResidentSequence(self.query, device)
        try:
            res = seq.align_windows(mat, del_, ext, starts=[0], lens=[len(self.query)], want_alignments=True,
                                    want_matrices=want_matrices, numeric=numeric)
        finally:
            seq.close()
        if res.status[0] != N.OK:
            _raise_status(int(res.status[0]))
        return AlignmentResult(res.H, res.D, res.alignment(0, self.symbols), None, res.numeric)


class HeuristicPWMAligner:
    """aligner-core/src/heuristic/mod.rs:81-141: PWM alignment -> frequency matrix of the alignment
    -> transform_matrix -> ... until f stops improving; returns the last result and its matrix.
    Every pass is one device call; the loop itself is the reference's."""

    def __init__(self, symbols, query):
        self.symbols = symbols
        self.query = query

    @classmethod
    def from_str_seqs(cls, query, _target="", symbols=DNA):
        return cls(symbols, symbols.str_to_vec(query))

    @classmethod
    def from_seqs(cls, query, _target=(), symbols=DNA):
        return cls(symbols, _as_codes(query).copy())

    def perform_alignment(self, del_, ext, matrix, heuristics=None, want_matrices=True, device=-1):
        if heuristics is None:
            raise Error("MissingArgument")      # heuristic/mod.rs:110-113
        h = heuristics
        transformed = transform_matrix(matrix, h.kd, h.r_squared, h.frequencies)  # :115-116
        max_f = 0.0
        while True:
            result = PWMAligner(self.symbols, self.query).perform_alignment(
                del_, ext, transformed, None, want_matrices=want_matrices, device=device)
            if result.alignment.f > max_f:      # :127-135
                max_f = result.alignment.f
                transformed = transform_matrix(result.alignment.get_frequency_matrix(), h.kd, h.r_squared,
                                               h.frequencies)
            else:                               # :136-139
                result.matrix = transformed
                return result
