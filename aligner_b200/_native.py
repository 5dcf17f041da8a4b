"""ctypes binding of include/aligner_b200.h (the C-ABI shared library).

The library is built in-tree by ``__graft_entry__.build()`` (``make -C aligner_b200/csrc``).
There is no CPU fallback: if the library is missing, loading raises; if no CUDA device is
usable, every compute call returns ``ALNR_CUDA_ERROR`` and the wrappers raise.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libaligner_b200.so")

OK = 0
STATUS_NAMES = {
    0: "OK", 1: "ProteinNotFound", 2: "CharIsNotMatchable", 3: "UnnecessaryArgument",
    4: "MissingArgument", 5: "ResultIsEmpty", 6: "CalculationError", 7: "ValidationError",
    8: "MatrixShapeError", 9: "WrongMatrixSpecified", 10: "EmptyInput", 11: "NoPositiveCell",
    12: "SymbolOutOfRange", 13: "CudaError", 14: "InvalidArgument", 15: "OutOfMemory",
}
GLOBAL, LOCAL = 0, 1
NUM_AUTO, NUM_I16X2, NUM_I32, NUM_F64 = 0, 1, 2, 3
FITNESS_BLOCK = 256

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
i32p = C.POINTER(C.c_int32)
f64p = C.POINTER(C.c_double)


class Params(C.Structure):
    _fields_ = [("mode", C.c_int), ("volume", C.c_int), ("del_", C.c_double), ("ext", C.c_double),
                ("matrix", f64p), ("device", C.c_int), ("numeric", C.c_int)]


class PairResult(C.Structure):
    _fields_ = [("f", C.c_double), ("corner_score", C.c_double), ("coords", C.c_uint64 * 4),
                ("q_aln", u8p), ("t_aln", u8p), ("aln_len", C.c_uint64), ("H", f64p), ("D", u8p),
                ("numeric", C.c_int)]


class Batch(C.Structure):
    _fields_ = [("q", u8p), ("q_off", u64p), ("t", u8p), ("t_off", u64p), ("count", C.c_uint64),
                ("q2", u32p), ("t2", u32p)]


class BatchOut(C.Structure):
    _fields_ = [("f", f64p), ("corner", f64p), ("coords", u64p), ("aln_off", u64p),
                ("aln_len", u32p), ("q_aln", u8p), ("t_aln", u8p), ("status", i32p),
                ("end_yx", u32p), ("moves_off", u64p), ("moves", u32p)]


class HeuristicItems(C.Structure):
    _fields_ = [("pair_idx", u32p), ("mat_idx", u32p), ("count", C.c_uint64), ("matrices", f64p),
                ("n_matrices", C.c_uint32), ("kd", C.c_double), ("r_squared", C.c_double),
                ("frequencies", f64p)]


class PwmParams(C.Structure):
    _fields_ = [("del_", C.c_double), ("ext", C.c_double), ("pwm", f64p), ("rows", C.c_uint32),
                ("length", C.c_uint32), ("device", C.c_int), ("numeric", C.c_int)]


class Windows(C.Structure):
    _fields_ = [("count", C.c_uint64), ("starts", u64p), ("lens", u32p), ("first", C.c_uint64),
                ("step", C.c_uint64), ("window", C.c_uint32)]


class PwmOut(C.Structure):
    _fields_ = [("f", f64p), ("coords", u64p), ("status", i32p), ("aln_off", u64p), ("aln_len", u32p),
                ("numbered", u32p), ("query", u8p), ("aln_slot", C.c_uint64), ("freq", f64p),
                ("H", f64p), ("D", u8p)]


class HeuristicOut(C.Structure):
    _fields_ = [("f", f64p), ("passes", i32p), ("status", i32p), ("coords", u64p), ("matrices", f64p)]


# every symbol include/aligner_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "alnr_strerror": (C.c_char_p, [C.c_int]),
    "alnr_last_error_detail": (C.c_char_p, []),
    "alnr_device_count": (C.c_int, []),
    "alnr_build_info": (C.c_char_p, []),
    "alnr_kernel_launch_count": (C.c_uint64, []),
    "alnr_align_pair": (C.c_int, [C.POINTER(Params), u8p, C.c_uint64, u8p, C.c_uint64, C.c_int,
                                  C.POINTER(PairResult)]),
    "alnr_free_pair_result": (None, [C.POINTER(PairResult)]),
    "alnr_align_batch": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), C.POINTER(BatchOut)]),
    "alnr_align_batch_multi": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), C.POINTER(BatchOut), C.c_int]),
    "alnr_matrix_search_multi": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), f64p, C.c_uint32, C.c_double,
                                           C.c_double, f64p, f64p, u32p, f64p, f64p, i32p, C.c_int]),
    "alnr_batch_create": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), C.POINTER(C.c_void_p)]),
    "alnr_batch_create_compact": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), C.POINTER(C.c_void_p)]),
    "alnr_moves_capacity": (C.c_uint64, [C.POINTER(Batch)]),
    "alnr_expand_alignments": (C.c_int, [C.c_int, C.c_int, C.POINTER(Batch), C.POINTER(BatchOut), C.c_uint64,
                                         C.c_uint64, C.POINTER(BatchOut), f64p, C.c_int]),
    "alnr_batch_run": (C.c_int, [C.c_void_p]),
    "alnr_batch_sync": (C.c_int, [C.c_void_p]),
    "alnr_batch_download": (C.c_int, [C.c_void_p, C.POINTER(BatchOut)]),
    "alnr_batch_stream": (C.c_void_p, [C.c_void_p]),
    "alnr_batch_run_timed": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "alnr_batch_numeric": (C.c_int, [C.c_void_p]),
    "alnr_batch_rerun_count": (C.c_uint64, [C.c_void_p]),
    "alnr_batch_destroy": (None, [C.c_void_p]),
    "alnr_score_batch": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), f64p, f64p, u32p, i32p]),
    "alnr_batch_create_scores": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), C.POINTER(C.c_void_p)]),
    "alnr_batch_download_scores": (C.c_int, [C.c_void_p, f64p, f64p, u32p, i32p]),
    "alnr_score_batch_noalloc": (C.c_int, [C.POINTER(Params), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p]),
    "alnr_noalloc_scratch_bytes": (C.c_uint64, []),
    "alnr_pwm_seq_create": (C.c_int, [u8p, C.c_uint64, C.c_int, C.POINTER(C.c_void_p)]),
    "alnr_pwm_seq_len": (C.c_uint64, [C.c_void_p]),
    "alnr_pwm_seq_destroy": (None, [C.c_void_p]),
    "alnr_pwm_align_windows": (C.c_int, [C.POINTER(PwmParams), C.c_void_p, C.POINTER(Windows),
                                         C.POINTER(PwmOut)]),
    "alnr_pwm_last_numeric": (C.c_int, []),
    "alnr_transform_matrix": (C.c_int, [f64p, C.c_int, C.c_int, C.c_double, C.c_double, f64p, f64p]),
    "alnr_heuristic_batch": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), C.POINTER(HeuristicItems),
                                       C.POINTER(HeuristicOut)]),
    "alnr_matrix_search": (C.c_int, [C.POINTER(Params), C.POINTER(Batch), f64p, C.c_uint32, C.c_double,
                                     C.c_double, f64p, f64p, u32p, f64p, f64p, i32p]),
    "alnr_fitness_block_sums": (C.c_int, [f64p, i32p, C.c_uint64, C.c_uint32, f64p]),
    "alnr_fitness_from_blocks": (C.c_int, [f64p, C.c_uint64, C.c_uint32, f64p]),
    "alnr_measure_int_peak": (C.c_int, [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "alnr_measure_fp64_peak": (C.c_int, [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "alnr_host_alloc": (C.c_void_p, [C.c_uint64]),
    "alnr_host_free": (None, [C.c_void_p]),
    "alnr_release_cached_memory": (None, []),
}

_lib = None


class AlignerError(RuntimeError):
    def __init__(self, status, detail=""):
        self.status = status
        self.name = STATUS_NAMES.get(status, "status %d" % status)
        super().__init__("%s%s" % (self.name, (": " + detail) if detail else ""))


def lib():
    """Load (once) and return the C-ABI library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(the CUDA extension is mandatory; there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(status):
    if status != OK:
        detail = lib().alnr_last_error_detail()
        raise AlignerError(status, detail.decode() if detail and status in (13, 14, 15) else "")
