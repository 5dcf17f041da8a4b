"""Compact wire formats (include/aligner_b200.h: alnr_batch.q2 / .t2, alnr_batch_out.moves).

CPU part: the host-side pieces — 2-bit packing and alnr_expand_alignments — against the oracle's
strings (the move string of an alignment is read off the oracle's own aligned strings).
GPU part: every traceback family in the compact result format, expanded, must equal the string
format of the same call; 2-bit inputs must equal byte inputs.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200 import workloads as W
from aligner_b200.batch import CompactResult, _ptr

M_DNA = ab.dna_matrix(5, -4)


def ragged_dna(count, lo, hi, seed=7):
    rng = np.random.default_rng(seed)
    qs, ts = [], []
    for _ in range(count):
        n = int(rng.integers(lo, hi + 1))
        q = rng.integers(0, 4, n, dtype=np.uint8)
        t = q.copy()
        sub = rng.random(n) < 0.15
        t[sub] = rng.integers(0, 4, int(sub.sum()), dtype=np.uint8)
        keep = rng.random(n) > 0.03
        t = t[keep]
        if len(t) == 0:
            t = q[:1].copy()
        qs.append(q)
        ts.append(t)
    return ab.PackedPairs.from_lists(qs, ts)


def moves_of(qa, ta):
    """2-bit move words of one alignment, read backwards from the column before the seed."""
    L = len(qa)
    codes = [0 if qa[p] == 98 else (1 if ta[p] == 98 else 2) for p in range(L - 2, -1, -1)]
    words = np.zeros((len(codes) + 15) // 16, dtype=np.uint32)
    for k, c in enumerate(codes):
        words[k >> 4] |= np.uint32(c << (2 * (k & 15)))
    return words


def compact_from_oracle(pairs, mode, ref):
    """A CompactResult filled from the oracle's outputs (what the device would return)."""
    cr = CompactResult(pairs, N.LOCAL if mode == "local" else N.GLOBAL, 4, want_corner=True)
    off = 0
    for i in range(pairs.count):
        L = int(ref["aln_len"][i])
        cr.aln_len[i] = L
        cr.status[i] = 0 if L else 11
        cr.moves_off[i] = off
        if L:
            co = ref["coords"][i]
            cr.end_yx[i] = (int(co[3]) - (1 if mode == "local" else 0), int(co[1]) - (1 if mode == "local" else 0))
            w = moves_of(ref["q_aln"][i, :L], ref["t_aln"][i, :L])
            cr.moves[off:off + len(w)] = w
            off += len(w)
    cr.f[:] = ref["f"]
    return cr


def test_pack2bit_layout():
    pairs = ragged_dna(40, 1, 70)
    p2 = pairs.pack2bit()
    woff = 0
    for i in range(pairs.count):
        q, _ = pairs.pair(i)
        for k, c in enumerate(q):
            assert (int(p2.q2[woff + k // 16]) >> (2 * (k % 16))) & 3 == c
        woff += (len(q) + 15) // 16
    assert woff == len(p2.q2)


@pytest.mark.parametrize("mode,d,e", [("local", 10.0, 2.0), ("global", 6.0, 6.0)])
@pytest.mark.parametrize("packed", [False, True])
def test_expand_matches_oracle_strings(mode, d, e, packed):
    pairs = ragged_dna(300, 20, 90, seed=11)
    ref = O.align_batch(mode, pairs, d, e, M_DNA, threads=4)
    cr = compact_from_oracle(pairs, mode, ref)
    if packed:
        cr.pairs = pairs.pack2bit()
    for first, count in ((0, pairs.count), (37, 101), (299, 1)):
        ex = cr.expand(first, count, want_freq=True, threads=3)
        for r in range(count):
            i = first + r
            L = int(ref["aln_len"][i])
            assert int(ex.aln_len[r]) == L
            qa, ta = ex.alignment(r)
            assert (qa == ref["q_aln"][i, :L]).all() and (ta == ref["t_aln"][i, :L]).all()
            assert (ex.coords[r] == ref["coords"][i]).all()
            fm = np.zeros((4, 4))
            O.lib().orc_frequency_matrix(_ptr(np.ascontiguousarray(qa), N.u8p), _ptr(np.ascontiguousarray(ta), N.u8p),
                                         L, 4, _ptr(fm, N.f64p))
            assert (ex.freq[r] == fm).all()
    # without moves_off the offsets are re-derived (first must be 0)
    cr.moves_off = None
    ex = cr.expand(0, pairs.count)
    i = pairs.count - 1
    L = int(ref["aln_len"][i])
    assert (ex.alignment(i)[0] == ref["q_aln"][i, :L]).all()


def test_expand_rejects_bad_arguments():
    pairs = ragged_dna(8, 10, 20)
    ref = O.align_batch("local", pairs, 10.0, 2.0, M_DNA)
    cr = compact_from_oracle(pairs, "local", ref)
    b, c = pairs.c_struct(), cr.c_struct()
    out = ab.BatchResult(pairs)
    e = out.c_struct()
    L = N.lib()
    assert L.alnr_expand_alignments(N.LOCAL, 4, C.byref(b), C.byref(c), 4, 9, C.byref(e), N.f64p(), 1) == 14
    cr.moves_off = None
    c = cr.c_struct()
    assert L.alnr_expand_alignments(N.LOCAL, 4, C.byref(b), C.byref(c), 1, 2, C.byref(e), N.f64p(), 1) == 14


# ---- GPU ---------------------------------------------------------------------------------------

def _same_as_strings(pairs, mode, d, e, numeric=N.NUM_AUTO, packed=False):
    src = pairs.pack2bit() if packed else pairs
    ref = ab.align_batch(mode, pairs, M_DNA, d, e, numeric=numeric)
    cr = ab.align_batch(mode, src, M_DNA, d, e, numeric=numeric, compact=True)
    assert (cr.f == ref.f).all() and (cr.corner == ref.corner).all() and (cr.status == ref.status).all()
    assert (cr.aln_len == ref.aln_len).all() and (cr.coords == ref.coords).all()
    words = (cr.aln_len.astype(np.int64) + 14) >> 4
    assert (cr.moves_off.astype(np.int64) == np.cumsum(words) - words).all()  # dense, pair order
    ex = cr.expand(threads=4)
    assert (ex.coords == ref.coords).all() and (ex.aln_len == ref.aln_len).all()
    for i in range(pairs.count):
        qa, ta = ex.alignment(i)
        qr, tr = ref.alignment(i)
        assert (qa == qr).all() and (ta == tr).all(), i
    ok = cr.status == 0
    if mode == N.LOCAL:
        assert (cr.end_yx[ok, 0].astype(np.uint64) + 1 == ref.coords[ok, 3]).all()
        assert (cr.end_yx[ok, 1].astype(np.uint64) + 1 == ref.coords[ok, 1]).all()
    return cr


@pytest.mark.gpu
@pytest.mark.parametrize("packed", [False, True])
@pytest.mark.parametrize("mode,d,e", [(N.LOCAL, 10, 2), (N.LOCAL, 6, 6), (N.GLOBAL, 8, 8)])
def test_compact_strip_kernel(mode, d, e, packed):
    """Uniform short reads on the packed strip kernel: move-string traceback chunk by chunk."""
    pairs = W.dna_read_batch(1000, 3000, n=150)
    _same_as_strings(pairs, mode, d, e, numeric=N.NUM_I16X2, packed=packed)


@pytest.mark.gpu
@pytest.mark.parametrize("packed", [False, True])
def test_compact_streamed_chunks(packed, monkeypatch):
    """Several chunks through the three-stream pipeline: the running total of move words."""
    monkeypatch.setenv("ALNR_STREAM_CHUNK", "4096")
    pairs = W.dna_read_batch(0, 20000, n=100)
    _same_as_strings(pairs, N.LOCAL, 10, 2, packed=packed)


@pytest.mark.gpu
@pytest.mark.parametrize("packed", [False, True])
@pytest.mark.parametrize("mode,d,e", [(N.LOCAL, 10, 2), (N.GLOBAL, 8, 8)])
def test_compact_ragged(mode, d, e, packed):
    """Ragged batch (sorted order, warp kernel, string traceback): moves derived from the strings."""
    pairs = ragged_dna(700, 5, 260, seed=3)
    _same_as_strings(pairs, mode, d, e, packed=packed)


@pytest.mark.gpu
def test_compact_small_and_long():
    _same_as_strings(W.dna_read_batch(5, 7, n=64), N.LOCAL, 10, 2)           # below 32 pairs
    q, t = W.dna_long_pair(6000, seed=9)
    _same_as_strings(ab.PackedPairs.from_lists([q], [t]), N.GLOBAL, 8, 8)    # tile pipeline + window walk
    _same_as_strings(ab.PackedPairs.from_lists([q], [t]), N.GLOBAL, 8, 8, packed=True)


@pytest.mark.gpu
def test_compact_device_batch_and_oracle():
    pairs = W.dna_read_batch(77, 2048, n=150)
    db = ab.DeviceBatch(N.LOCAL, pairs.pack2bit(), M_DNA, 10, 2, numeric=N.NUM_I16X2, compact=True)
    db.run()
    db.run()
    cr = db.download()
    db.close()
    ref = O.align_batch("local", pairs, 10.0, 2.0, M_DNA, threads=4)
    ex = cr.expand()
    assert (cr.f == ref["f"]).all() and (ex.coords == ref["coords"]).all()
    for i in range(pairs.count):
        L = int(ref["aln_len"][i])
        qa, ta = ex.alignment(i)
        assert (qa == ref["q_aln"][i, :L]).all() and (ta == ref["t_aln"][i, :L]).all()


@pytest.mark.gpu
def test_packed_input_needs_dna_and_strings_stay_default():
    pairs = W.dna_read_batch(0, 64, n=50)
    with pytest.raises(ab.AlignerError):
        ab.align_batch(N.LOCAL, pairs.pack2bit(), ab.get_blosum62(), 10, 2)
    res = ab.align_batch(N.LOCAL, pairs.pack2bit(), M_DNA, 10, 2)  # packed in, strings out
    ref = ab.align_batch(N.LOCAL, pairs, M_DNA, 10, 2)
    assert (res.f == ref.f).all() and (res.q_aln == ref.q_aln).all() and (res.t_aln == ref.t_aln).all()
    sc = ab.score_batch(N.LOCAL, pairs.pack2bit(), M_DNA, 10, 2)
    assert (sc.f == ref.f).all()
