"""Deterministic synthetic inputs for the BASELINE.json configs.

Counter-based generator: every symbol is a pure function of (seed, pair index, position),
built on the splitmix64 finaliser, so any sub-range of a workload can be regenerated
without the rest (parity subsamples, per-rank shards).
"""
import numpy as np

from .batch import PackedPairs

_G = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def _mix64(z):
    z = (z ^ (z >> np.uint64(30))) * _M1
    z = (z ^ (z >> np.uint64(27))) * _M2
    return z ^ (z >> np.uint64(31))


def _draw(seed, pair_idx, pos, stream):
    """uint64 hash per (pair, position, stream); arrays broadcast."""
    with np.errstate(over="ignore"):
        s = (np.uint64(seed) + pair_idx.astype(np.uint64)) * _G
        return _mix64(s + pos.astype(np.uint64) * _M1 + np.uint64(stream) * _M2)


def dna_read_pairs(first, count, n=150, seed=0xA11C0DE, sub_rate=0.10, block=1 << 18):
    """C3: `count` DNA pairs starting at pair index `first`; query uniform i.i.d. over
    A,T,C,G codes; target = query with each base independently replaced (prob sub_rate)
    by a uniform base.  Returns (q2d, t2d) uint8 arrays of shape (count, n)."""
    q = np.empty((count, n), dtype=np.uint8)
    t = np.empty((count, n), dtype=np.uint8)
    pos = np.arange(n, dtype=np.uint64)[None, :]
    thr = np.uint64(int(sub_rate * 65536.0))
    for b0 in range(0, count, block):
        b1 = min(b0 + block, count)
        idx = np.arange(first + b0, first + b1, dtype=np.uint64)[:, None]
        h = _draw(seed, idx, pos, 1)
        qb = (h & np.uint64(3)).astype(np.uint8)
        g = _draw(seed, idx, pos, 2)
        replace = (g & np.uint64(0xFFFF)) < thr
        rb = ((g >> np.uint64(16)) & np.uint64(3)).astype(np.uint8)
        q[b0:b1] = qb
        t[b0:b1] = np.where(replace, rb, qb)
    return q, t


def dna_read_batch(first, count, n=150, seed=0xA11C0DE, sub_rate=0.10):
    q, t = dna_read_pairs(first, count, n, seed, sub_rate)
    return PackedPairs.uniform(q, t)


def dna_long_pair(n=100000, seed=4, sub_rate=0.15, indel_rate=0.01):
    """C4 synthetic: one long DNA pair; target = query with substitutions and
    single-base insertions/deletions."""
    pos = np.arange(n, dtype=np.uint64)
    zero = np.zeros(1, dtype=np.uint64)
    q = (_draw(seed, zero, pos, 1) & np.uint64(3)).astype(np.uint8)
    g = _draw(seed, zero, pos, 2)
    u = (g & np.uint64(0xFFFF)).astype(np.float64) / 65536.0
    rb = ((g >> np.uint64(16)) & np.uint64(3)).astype(np.uint8)
    ins = ((g >> np.uint64(20)) & np.uint64(3)).astype(np.uint8)
    out = []
    # vectorised edit script: op 0 keep/sub, 1 delete, 2 insert-after
    op = np.zeros(n, dtype=np.uint8)
    op[u < indel_rate / 2] = 1
    op[(u >= indel_rate / 2) & (u < indel_rate)] = 2
    base = np.where((u >= indel_rate) & (u < indel_rate + sub_rate), rb, q)
    keep = op != 1
    reps = np.where(op == 2, 2, 1)[keep]
    tb = np.repeat(base[keep], reps)
    # second copy of a repeated position is the inserted base
    ends = np.cumsum(reps) - 1
    two = reps == 2
    tb[ends[two]] = ins[keep][two]
    del out
    return q, tb.astype(np.uint8)


# Robinson & Robinson style background frequencies for the 20 standard residues, in the
# reference's Protein code order A R N D C Q E G H I L K M F P S T W Y V.
_AA_FREQ = np.array([0.078, 0.051, 0.045, 0.054, 0.019, 0.043, 0.063, 0.074, 0.022, 0.051,
                     0.091, 0.057, 0.022, 0.039, 0.052, 0.071, 0.058, 0.013, 0.032, 0.065])


def protein_pairs(first, count, len_lo=270, len_hi=330, seed=5, sub_rate=0.30):
    """C5: protein pairs, lengths uniform in [len_lo, len_hi], residues 0..19 drawn from
    background frequencies; target = query with sub_rate substitutions."""
    cdf = np.cumsum(_AA_FREQ / _AA_FREQ.sum())
    qs, ts = [], []
    for i in range(first, first + count):
        idx = np.array([i], dtype=np.uint64)
        L = len_lo + int(_draw(seed, idx, np.zeros(1, dtype=np.uint64), 7)[0] % np.uint64(len_hi - len_lo + 1))
        pos = np.arange(L, dtype=np.uint64)
        u = (_draw(seed, idx, pos, 1) >> np.uint64(11)).astype(np.float64) / float(1 << 53)
        q = np.minimum(np.searchsorted(cdf, u), 19).astype(np.uint8)
        g = _draw(seed, idx, pos, 2)
        u2 = (_draw(seed, idx, pos, 3) >> np.uint64(11)).astype(np.float64) / float(1 << 53)
        r = np.minimum(np.searchsorted(cdf, u2), 19).astype(np.uint8)
        rep = (g & np.uint64(0xFFFF)).astype(np.float64) / 65536.0 < sub_rate
        qs.append(q)
        ts.append(np.where(rep, r, q).astype(np.uint8))
    return PackedPairs.from_lists(qs, ts)


def candidate_matrices(k, dim=24, seed=5, threshold=24.6):
    """C5 candidates by the rule of get_population (aligner-web/src/repository/mod.rs:352-377):
    uniform {-1,0,1}^(dim x dim), rejected when the L2 distance to an accepted matrix is below
    get_threshold(dim) (aligner-helpers/src/matrices/mod.rs:8-17).  The reference draws from an
    unseeded thread_rng, so only the rule is reproducible, not its matrices."""
    out = np.zeros((0, dim, dim), dtype=np.float64)
    trial = 0
    pos = np.arange(dim * dim, dtype=np.uint64)
    while out.shape[0] < k:
        h = _draw(seed, np.array([trial], dtype=np.uint64), pos, 11)
        trial += 1
        cand = ((h >> np.uint64(20)) % np.uint64(3)).astype(np.float64).reshape(dim, dim) - 1.0
        if out.shape[0]:
            d = np.sqrt(((out - cand[None]) ** 2).sum(axis=(1, 2)))
            if (d < threshold).any():
                continue
        out = np.concatenate([out, cand[None]], axis=0)
    return out


def pair_frequencies(pairs, V):
    """Per-pair residue frequencies over both sequences, symbol order (the intent of
    aligner-web/src/bin/dispatcher/handlers.rs:209-222)."""
    fr = np.zeros((pairs.count, V), dtype=np.float64)
    for i in range(pairs.count):
        q, t = pairs.pair(i)
        c = np.bincount(np.concatenate([q, t]), minlength=V).astype(np.float64)
        fr[i] = c / c.sum()
    return fr


def dna_read_pairs_torch(first, count, n=150, seed=0xA11C0DE, sub_rate=0.10, device="cuda", block=1 << 20,
                         out_q=None, out_t=None, out_q2=None, out_t2=None):
    """Same pairs as dna_read_pairs, generated on the GPU with torch (bench input setup only —
    not the product path).  int64 arithmetic wraps like uint64; logical shifts are emulated.
    Returns (q2d, t2d) uint8 host arrays (written into out_q/out_t when given).  out_q2 / out_t2:
    (count, ceil(n/16)) uint32 host arrays that receive the same pairs in the compact input format
    (2-bit codes, include/aligner_b200.h alnr_batch.q2)."""
    import torch

    def i64(v):
        v &= (1 << 64) - 1
        return v - (1 << 64) if v >= (1 << 63) else v

    G, M1, M2 = i64(0x9E3779B97F4A7C15), i64(0xBF58476D1CE4E5B9), i64(0x94D049BB133111EB)

    def lsr(z, k):
        return (z >> k) & ((1 << (64 - k)) - 1)

    def mix(z):
        z = (z ^ lsr(z, 30)) * M1
        z = (z ^ lsr(z, 27)) * M2
        return z ^ lsr(z, 31)

    q = out_q if out_q is not None else np.empty((count, n), dtype=np.uint8)
    t = out_t if out_t is not None else np.empty((count, n), dtype=np.uint8)
    pos = torch.arange(n, dtype=torch.int64, device=device)[None, :] * M1
    thr = int(sub_rate * 65536.0)
    nw = (n + 15) // 16
    sh = (2 * torch.arange(16, dtype=torch.int64, device=device))[None, None, :]

    def pack(codes):  # (B, n) uint8 -> (B, nw) uint32 bit patterns as int32
        padded = torch.zeros((codes.shape[0], nw * 16), dtype=torch.int64, device=device)
        padded[:, :n] = codes
        w = (padded.view(-1, nw, 16) << sh).sum(dim=2)
        return w.to(torch.int32).cpu().numpy().view(np.uint32)

    for b0 in range(0, count, block):
        b1 = min(b0 + block, count)
        idx = torch.arange(first + b0, first + b1, dtype=torch.int64, device=device)[:, None]
        s = (idx + i64(seed)) * G + pos
        h = mix(s + i64(1 * 0x94D049BB133111EB))
        qb = (h & 3).to(torch.uint8)
        g = mix(s + i64(2 * 0x94D049BB133111EB))
        replace = (g & 0xFFFF) < thr
        rb = (lsr(g, 16) & 3).to(torch.uint8)
        tb = torch.where(replace, rb, qb)
        q[b0:b1] = qb.cpu().numpy()
        t[b0:b1] = tb.cpu().numpy()
        if out_q2 is not None:
            out_q2[b0:b1] = pack(qb)
            out_t2[b0:b1] = pack(tb)
    return q, t


def dna_chromosome_with_repeats(length, repeat_length=300, copies=40, seed=7, sub_rate=0.25, indel_rate=0.02):
    """N1 synthetic (the shape of bin/latent-repeat-search's input): an i.i.d. DNA sequence with
    `copies` diverged copies of one random repeat planted at random places.  Returns (codes uint8,
    repeat uint8, planted start positions)."""
    rng = np.random.default_rng(seed)
    seq = rng.integers(0, 4, length).astype(np.uint8)
    repeat = rng.integers(0, 4, repeat_length).astype(np.uint8)
    starts = np.sort(rng.choice(max(length - 2 * repeat_length, 1), size=copies, replace=False))
    for s in starts:
        copy = repeat.copy()
        mut = rng.random(repeat_length) < sub_rate
        copy[mut] = rng.integers(0, 4, int(mut.sum()))
        keep = rng.random(repeat_length) >= indel_rate
        copy = copy[keep]
        seq[s:s + len(copy)] = copy
    return seq, repeat, starts


def pwm_from_repeat(repeat, match=1.0, mismatch=-1.0):
    """4 x L profile scoring `match` for the repeat's base at each position, `mismatch` otherwise."""
    L = len(repeat)
    pwm = np.full((4, L), float(mismatch))
    pwm[repeat.astype(np.intp), np.arange(L)] = float(match)
    return pwm
