"""ctypes access to oracle/liboracle.so — the CHECKER used by tests, smoke() and the
cpu_baseline leg of bench.py.  Never imported by aligner_b200/."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
f64p = C.POINTER(C.c_double)
i32p = C.POINTER(C.c_int)

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        L = C.CDLL(LIB)
        L.orc_align_pair.restype = C.c_int
        L.orc_align_pair.argtypes = [C.c_int, u8p, C.c_size_t, u8p, C.c_size_t, C.c_double, C.c_double,
                                     f64p, C.c_int, f64p, u8p, u8p, u8p, C.c_size_t,
                                     C.POINTER(C.c_size_t), u64p, f64p, f64p]
        L.orc_align_pair_lean.restype = C.c_int
        L.orc_align_pair_lean.argtypes = [C.c_int, u8p, C.c_size_t, u8p, C.c_size_t, C.c_double,
                                          C.c_double, f64p, C.c_int, u8p, u8p, C.c_size_t,
                                          C.POINTER(C.c_size_t), u64p, f64p, f64p, C.c_char_p]
        L.orc_transform_matrix.restype = C.c_int
        L.orc_transform_matrix.argtypes = [f64p, C.c_int, C.c_int, C.c_double, C.c_double, f64p, f64p]
        L.orc_frequency_matrix.restype = None
        L.orc_frequency_matrix.argtypes = [u8p, u8p, C.c_size_t, C.c_int, f64p]
        L.orc_get_alignment.restype = None
        L.orc_get_alignment.argtypes = [u8p, u8p, C.c_size_t, f64p, C.c_int, u8p]
        L.orc_heuristic.restype = C.c_int
        L.orc_heuristic.argtypes = [u8p, C.c_size_t, u8p, C.c_size_t, C.c_double, C.c_double, f64p,
                                    C.c_int, C.c_double, C.c_double, f64p, u8p, u8p, C.c_size_t,
                                    C.POINTER(C.c_size_t), u64p, f64p, f64p, i32p]
        L.orc_align_batch.restype = C.c_int
        L.orc_align_batch.argtypes = [C.c_int, u8p, u64p, u8p, u64p, C.c_size_t, C.c_double,
                                      C.c_double, f64p, C.c_int, C.c_int, f64p, f64p, u64p, u64p,
                                      u8p, u8p, C.c_size_t, i32p]
        L.orc_heuristic_batch.restype = C.c_int
        L.orc_heuristic_batch.argtypes = [u8p, u64p, u8p, u64p, u32p, u32p, C.c_size_t, C.c_double,
                                          C.c_double, f64p, C.c_int, C.c_double, C.c_double, f64p,
                                          C.c_int, f64p, i32p, f64p, i32p]
        L.orc_pwm_align.restype = C.c_int
        L.orc_pwm_align.argtypes = [u8p, C.c_size_t, f64p, C.c_size_t, C.c_size_t, C.c_double, C.c_double,
                                    f64p, u8p, u32p, u8p, C.c_size_t, C.POINTER(C.c_size_t), u64p, f64p]
        L.orc_pwm_frequency_matrix.restype = None
        L.orc_pwm_frequency_matrix.argtypes = [u32p, u8p, C.c_size_t, C.c_size_t, f64p]
        L.orc_pwm_windows.restype = None
        L.orc_pwm_windows.argtypes = [u8p, u64p, u32p, C.c_size_t, f64p, C.c_size_t, C.c_double, C.c_double,
                                      C.c_int, f64p, u64p, u64p, u32p, u8p, C.c_size_t, i32p]
        L.orc_ndarray_sum.restype = C.c_double
        L.orc_ndarray_sum.argtypes = [f64p, C.c_size_t]
        L.orc_find_roots_quadratic.restype = C.c_int
        L.orc_find_roots_quadratic.argtypes = [C.c_double, C.c_double, C.c_double, f64p]
        L.orc_get_threshold.restype = C.c_double
        L.orc_get_threshold.argtypes = [C.c_int]
        _lib = L
    return _lib


def _p(a, ty):
    return a.ctypes.data_as(ty) if a is not None else ty()


MODES = {"global": 0, "local": 1, 0: 0, 1: 1}
STATUS_TO_ALNR = {0: 0, 1: 10, 2: 11, 3: 12, 4: 9, 5: 14, 6: 8}  # oracle status -> alnr_status


def align_pair(mode, q, t, del_, ext, matrix, want_matrices=True, lean=False):
    """Returns dict(status, f, corner, coords, q_aln, t_aln, H, D[, d_sha])."""
    q = np.ascontiguousarray(q, dtype=np.uint8)
    t = np.ascontiguousarray(t, dtype=np.uint8)
    mat = np.ascontiguousarray(matrix, dtype=np.float64)
    V = mat.shape[0]
    n, m = len(q), len(t)
    cap = n + m + 2
    qa = np.zeros(cap, dtype=np.uint8)
    ta = np.zeros(cap, dtype=np.uint8)
    L = C.c_size_t(0)
    coords = np.zeros(4, dtype=np.uint64)
    f = C.c_double(0)
    corner = C.c_double(0)
    out = {}
    if lean:
        sha = C.create_string_buffer(17)
        st = lib().orc_align_pair_lean(MODES[mode], _p(q, u8p), n, _p(t, u8p), m, del_, ext,
                                       _p(mat, f64p), V, _p(qa, u8p), _p(ta, u8p), cap, C.byref(L),
                                       _p(coords, u64p), C.byref(f), C.byref(corner), sha)
        out["d_sha"] = sha.value.decode()
        H = D = None
    else:
        H = np.zeros((m + 1, n + 1), dtype=np.float64) if want_matrices else None
        D = np.zeros((m + 1, n + 1), dtype=np.uint8) if want_matrices else None
        st = lib().orc_align_pair(MODES[mode], _p(q, u8p), n, _p(t, u8p), m, del_, ext, _p(mat, f64p),
                                  V, _p(H, f64p), _p(D, u8p), _p(qa, u8p), _p(ta, u8p), cap,
                                  C.byref(L), _p(coords, u64p), C.byref(f), C.byref(corner))
    out.update(status=st, f=f.value, corner=corner.value,
               coords=((int(coords[0]), int(coords[1])), (int(coords[2]), int(coords[3]))),
               q_aln=qa[:L.value].copy(), t_aln=ta[:L.value].copy(), H=H, D=D)
    return out


def align_batch(mode, pairs, del_, ext, matrix, threads=1, want_alignments=True):
    """pairs: aligner_b200.PackedPairs-like (q, q_off, t, t_off, count). Left-aligned slots."""
    mat = np.ascontiguousarray(matrix, dtype=np.float64)
    c = pairs.count
    n = np.diff(pairs.q_off).astype(np.int64)
    m = np.diff(pairs.t_off).astype(np.int64)
    slot = int((n + m).max()) + 2 if c else 2
    f = np.zeros(c)
    corner = np.zeros(c)
    coords = np.zeros((c, 4), dtype=np.uint64)
    alen = np.zeros(c, dtype=np.uint64)
    status = np.zeros(c, dtype=np.int32)
    qa = np.zeros((c, slot), dtype=np.uint8) if want_alignments else None
    ta = np.zeros((c, slot), dtype=np.uint8) if want_alignments else None
    lib().orc_align_batch(MODES[mode], _p(pairs.q, u8p), _p(pairs.q_off, u64p), _p(pairs.t, u8p),
                          _p(pairs.t_off, u64p), c, del_, ext, _p(mat, f64p), mat.shape[0], threads,
                          _p(f, f64p), _p(corner, f64p), _p(coords, u64p), _p(alen, u64p),
                          _p(qa, u8p), _p(ta, u8p), slot, _p(status, i32p))
    return dict(f=f, corner=corner, coords=coords, aln_len=alen, status=status, q_aln=qa, t_aln=ta)


def transform_matrix(matrix, kd, r_squared, freq):
    mat = np.ascontiguousarray(matrix, dtype=np.float64)
    fr = np.ascontiguousarray(freq, dtype=np.float64)
    out = np.zeros_like(mat)
    st = lib().orc_transform_matrix(_p(mat, f64p), mat.shape[0], mat.shape[1], kd, r_squared,
                                    _p(fr, f64p), _p(out, f64p))
    return st, out


def heuristic(q, t, del_, ext, matrix, kd, r_squared, freq):
    q = np.ascontiguousarray(q, dtype=np.uint8)
    t = np.ascontiguousarray(t, dtype=np.uint8)
    mat = np.ascontiguousarray(matrix, dtype=np.float64)
    fr = np.ascontiguousarray(freq, dtype=np.float64)
    V = mat.shape[0]
    cap = len(q) + len(t) + 2
    qa = np.zeros(cap, dtype=np.uint8)
    ta = np.zeros(cap, dtype=np.uint8)
    L = C.c_size_t(0)
    coords = np.zeros(4, dtype=np.uint64)
    f = C.c_double(0)
    om = np.zeros_like(mat)
    passes = C.c_int(0)
    st = lib().orc_heuristic(_p(q, u8p), len(q), _p(t, u8p), len(t), del_, ext, _p(mat, f64p), V, kd,
                             r_squared, _p(fr, f64p), _p(qa, u8p), _p(ta, u8p), cap, C.byref(L),
                             _p(coords, u64p), C.byref(f), _p(om, f64p), C.byref(passes))
    return dict(status=st, f=f.value, q_aln=qa[:L.value].copy(), t_aln=ta[:L.value].copy(),
                coords=((int(coords[0]), int(coords[1])), (int(coords[2]), int(coords[3]))),
                matrix=om, passes=passes.value)


def pwm_align(query, pwm, del_, ext, want_matrices=True):
    """orc_pwm_align -> dict(status, f, coords, numbered, q_aln, H, D)."""
    q = np.ascontiguousarray(query, dtype=np.uint8)
    mat = np.ascontiguousarray(pwm, dtype=np.float64)
    rows, L = mat.shape
    m = len(q)
    cap = m + L + 1
    num = np.zeros(cap, dtype=np.uint32)
    qa = np.zeros(cap, dtype=np.uint8)
    n = C.c_size_t(0)
    coords = np.zeros(4, dtype=np.uint64)
    f = C.c_double(0)
    H = np.zeros((m + 1, L + 1), dtype=np.float64) if want_matrices else None
    D = np.zeros((m + 1, L + 1), dtype=np.uint8) if want_matrices else None
    st = lib().orc_pwm_align(_p(q, u8p), m, _p(mat, f64p), rows, L, del_, ext, _p(H, f64p), _p(D, u8p),
                             _p(num, u32p), _p(qa, u8p), cap, C.byref(n), _p(coords, u64p), C.byref(f))
    return dict(status=st, f=f.value, numbered=num[:n.value].copy(), q_aln=qa[:n.value].copy(), H=H, D=D,
                coords=((int(coords[0]), int(coords[1])), (int(coords[2]), int(coords[3]))))


def pwm_frequency_matrix(numbered, q_aln, L):
    num = np.ascontiguousarray(numbered, dtype=np.uint32)
    qa = np.ascontiguousarray(q_aln, dtype=np.uint8)
    out = np.zeros((4, L), dtype=np.float64)
    lib().orc_pwm_frequency_matrix(_p(num, u32p), _p(qa, u8p), len(num), L, _p(out, f64p))
    return out


def pwm_windows(seq, starts, lens, pwm, del_, ext, threads=8, want_alignments=True):
    """orc_pwm_windows -> dict(f, coords, aln_len, numbered[count, slot], q_aln[count, slot], status)."""
    s = np.ascontiguousarray(seq, dtype=np.uint8)
    st_a = np.ascontiguousarray(starts, dtype=np.uint64)
    ln_a = np.ascontiguousarray(lens, dtype=np.uint32)
    mat = np.ascontiguousarray(pwm, dtype=np.float64)
    L = mat.shape[1]
    c = len(st_a)
    slot = L + (int(ln_a.max()) if c else 0) + 1
    f = np.zeros(c)
    coords = np.zeros((c, 4), dtype=np.uint64)
    alen = np.zeros(c, dtype=np.uint64)
    status = np.zeros(c, dtype=np.int32)
    num = np.zeros((c, slot), dtype=np.uint32) if want_alignments else None
    qa = np.zeros((c, slot), dtype=np.uint8) if want_alignments else None
    lib().orc_pwm_windows(_p(s, u8p), _p(st_a, u64p), _p(ln_a, u32p), c, _p(mat, f64p), L, del_, ext, threads,
                          _p(f, f64p), _p(coords, u64p), _p(alen, u64p), _p(num, u32p), _p(qa, u8p), slot,
                          _p(status, i32p))
    return dict(f=f, coords=coords, aln_len=alen, numbered=num, q_aln=qa, status=status)
