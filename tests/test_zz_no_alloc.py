"""alnr_score_batch_noalloc (aligner_b200/csrc/noalloc.cu): the score-only call that allocates nothing —
its scratch is a static device array — against the existing entry point alnr_score_batch, through
the C ABI, on two calls in a row (the shared scratch keeps the first call's contents).

No counterpart in the reference.  This file is named to
be collected LAST.  torch only provides the caller's device buffers."""
import ctypes as C

import numpy as np
import pytest

import aligner_b200 as ab
from aligner_b200 import _native as N
from aligner_b200.batch import make_params

pytestmark = pytest.mark.gpu


def _ragged(rng, count, lo, hi, V):
    qs, ts = [], []
    for _ in range(count):
        n = int(rng.integers(lo, hi + 1))
        q = rng.integers(0, V, n).astype(np.uint8)
        t = q.copy()
        mut = rng.random(n) < 0.2
        t[mut] = rng.integers(0, V, int(mut.sum()))
        t = t[int(rng.integers(0, 4)):]
        qs.append(q)
        ts.append(t if len(t) else q[:1].copy())
    return ab.PackedPairs.from_lists(qs, ts)


def _noalloc(mode, pairs, M, d, e, n_cap=None, m_cap=None, numeric=N.NUM_AUTO):
    """One call of the new entry point: device buffers from torch, outputs pre-filled with garbage."""
    import torch
    dev = torch.device("cuda", 0)
    up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).view(dt)).to(dev)
    q, t = up(pairs.q, np.uint8), up(pairs.t, np.uint8)
    qo, to = up(pairs.q_off, np.int64), up(pairs.t_off, np.int64)
    c = pairs.count
    f = torch.full((c,), -7.5, dtype=torch.float64, device=dev)
    corner = torch.full((c,), -7.5, dtype=torch.float64, device=dev)
    yx = torch.full((c, 2), 12345, dtype=torch.int32, device=dev)
    st = torch.full((c,), 99, dtype=torch.int32, device=dev)
    n_cap = int(np.diff(pairs.q_off).max()) if n_cap is None else n_cap
    m_cap = int(np.diff(pairs.t_off).max()) if m_cap is None else m_cap
    p = make_params(mode, M, d, e, 0, numeric)
    torch.cuda.synchronize()
    rc = N.lib().alnr_score_batch_noalloc(C.byref(p), q.data_ptr(), qo.data_ptr(), t.data_ptr(), to.data_ptr(), c,
                                          n_cap, m_cap, f.data_ptr(), corner.data_ptr(), yx.data_ptr(),
                                          st.data_ptr(), None)
    torch.cuda.synchronize()
    return rc, f.cpu().numpy(), corner.cpu().numpy(), yx.cpu().numpy().view(np.uint32), st.cpu().numpy()


def _compare(mode, pairs, M, d, e, numeric=N.NUM_AUTO):
    ref = ab.score_batch(mode, pairs, M, d, e, numeric=numeric)
    rc, f, corner, yx, st = _noalloc(mode, pairs, M, d, e, numeric=numeric)
    assert rc == 0
    assert (st == ref.status).all()
    ok = ref.status == 0
    # the tolerance of every parity test in this repository: bit-exact
    assert (f[ok] == ref.f[ok]).all() and (corner[ok] == ref.corner[ok]).all()
    assert (yx[ok] == ref.end_yx[ok]).all()
    return st


def test_noalloc_matches_score_batch_two_calls_in_a_row():
    rng = np.random.default_rng(99)
    dna = _ragged(rng, 700, 20, 180, 4)
    prot = _ragged(rng, 300, 30, 260, 24)
    Md, Mp = ab.dna_matrix(5, -4), ab.get_blosum62()
    # call 1 and call 2 share the static scratch: different shapes, modes and numeric domains
    _compare(N.LOCAL, dna, Md, 10, 2)
    _compare(N.GLOBAL, prot, Mp, 11, 1)
    _compare(N.LOCAL, prot, Mp + 0.125, 6.5, 1.25)         # non-integral: FP64 instantiation
    _compare(N.LOCAL, dna, Md, 10, 2)                        # the first problem again, over stale scratch
    _compare(N.GLOBAL, dna, Md, 8, 8, numeric=N.NUM_F64)


def test_noalloc_statuses_and_limits():
    rng = np.random.default_rng(5)
    qs = [rng.integers(0, 4, 40).astype(np.uint8) for _ in range(6)]
    ts = [q.copy() for q in qs]
    ts[1] = np.zeros(0, dtype=np.uint8)                      # empty target
    qs[2] = qs[2].copy(); qs[2][7] = 9                       # symbol outside the alphabet
    qs[3] = np.zeros(40, dtype=np.uint8)                     # all A against all T: no positive cell
    ts[3] = np.ones(40, dtype=np.uint8)
    pairs = ab.PackedPairs.from_lists(qs, ts)
    st = _compare(N.LOCAL, pairs, ab.dna_matrix(1, -1), 2, 1)
    assert st.tolist() == [0, 10, 12, 11, 0, 0]
    # a pair longer than promised is rejected per pair, never computed
    rc, f, corner, yx, st = _noalloc(N.LOCAL, pairs, ab.dna_matrix(1, -1), 2, 1, n_cap=40, m_cap=39)
    assert rc == 0 and st.tolist() == [14, 10, 14, 14, 14, 14]
    # beyond the static scratch: refused with the arithmetic in the detail text
    assert N.lib().alnr_noalloc_scratch_bytes() == 16 << 20
    rc, *_ = _noalloc(N.LOCAL, pairs, ab.dna_matrix(1, -1), 2, 1, m_cap=200000)
    assert rc == 14 and b"static scratch" in N.lib().alnr_last_error_detail()
    big = _ragged(rng, 9000, 450, 500, 4)
    rc, *_ = _noalloc(N.LOCAL, big, ab.dna_matrix(5, -4), 10, 2)
    assert rc == 14
