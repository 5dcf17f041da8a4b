"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/aligner_b200.h declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT

import aligner_b200 as ab
from aligner_b200 import _native as N


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "aligner_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(alnr_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    L = C.CDLL(N.LIB_PATH)
    syms = _declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(L, s), s
    # and the ctypes binding covers the same set
    assert set(syms) == set(N.SIGNATURES)


def test_strerror_and_build_info():
    L = N.lib()
    assert L.alnr_strerror(0) == b"ok"
    assert b"UnnecessaryArgument" in L.alnr_strerror(3)
    assert b"sm_100a" in L.alnr_build_info()


def test_no_cpu_fallback_without_gpu():
    if N.lib().alnr_device_count() > 0:
        pytest.skip("a GPU is present")
    a = ab.SimpleLocalAligner.from_str_seqs("HEAGAWGHEE", "PAWHEAE")
    with pytest.raises(ab.AlignerError) as ei:
        a.perform_alignment(11, 2, ab.get_blosum62())
    assert ei.value.status == 13


def test_reference_api_error_behaviour():
    # simple/mod.rs:49-51: Some(heuristics) -> Error::UnnecessaryArgument
    a = ab.SimpleGlobalAligner.from_str_seqs("HEAG", "PAW")
    with pytest.raises(ab.Error) as ei:
        a.perform_alignment(8, 8, ab.get_blosum62(), heuristics=ab.Heuristics(0, 0, np.ones(24) / 24))
    assert ei.value.kind == "UnnecessaryArgument"
    # enums.rs:230: unknown character -> CharIsNotMatchable
    with pytest.raises(ab.Error) as ei:
        ab.SimpleLocalAligner.from_str_seqs("HEAG!", "PAW")
    assert ei.value.kind == "CharIsNotMatchable"
    with pytest.raises(ab.Error):
        ab.DNA.str_to_vec("ACGT ")  # str_to_vec rejects the stray space ...
    assert list(ab.DNA.from_u8_vec(b"ACGT ")) == [0, 2, 3, 1]  # ... from_u8_vec skips it


def test_symbol_codes():
    assert ab.Protein.volume() == 24 and ab.DNA.volume() == 4
    assert ab.Protein.vec_to_str(ab.Protein.str_to_vec("ARNDCQEGHILKMFPSTWYVBJZX_+")) == "ARNDCQEGHILKMFPSTWYVBJZX_+"
    assert list(ab.DNA.str_to_vec("ATCG")) == [0, 1, 2, 3]
    assert ab.Protein.blank() == 98 and ab.Protein.pos() == 99
