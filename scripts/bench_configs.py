#!/usr/bin/env python
"""Side benchmarks for the BASELINE configs that are not the bench.py headline (C1, C2, C4, C5).

  python scripts/bench_configs.py [c4] [c4full] [c5] [c12]

Prints one JSON line per measurement (GCUPS, device time through the C-ABI call incl. copies)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import aligner_b200 as ab  # noqa: E402
from aligner_b200 import _native as N  # noqa: E402
from aligner_b200 import workloads as W  # noqa: E402
from aligner_b200.batch import PackedPairs  # noqa: E402


def timed_pair(mode, q, t, d, e, M, symbols, reps=3):
    cls = ab.SimpleGlobalAligner if mode == "global" else ab.SimpleLocalAligner
    al = cls.from_seqs(q, t, symbols=symbols)
    r = al.perform_alignment(d, e, M, want_matrices=False)  # warm-up
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        r = al.perform_alignment(d, e, M, want_matrices=False)
        best = min(best, time.perf_counter() - t0)
    return r, best


def emit(name, cells, secs, **kw):
    print(json.dumps(dict(config=name, cells=cells, seconds=secs, gcups=cells / secs / 1e9, **kw)), flush=True)


def c4(n, check):
    q, t = W.dna_long_pair(n, seed=4)
    M = ab.dna_matrix(5, -4)
    r, secs = timed_pair("global", q, t, 8, 8, M, ab.DNA)
    extra = {}
    if check:
        import oracle_lib as O
        t0 = time.perf_counter()
        ref = O.align_pair("global", q, t, 8.0, 8.0, M, lean=True)
        cpu = time.perf_counter() - t0
        ok = bool(r.corner_score == ref["corner"] and (r.alignment.query == ref["q_aln"]).all()
                  and (r.alignment.target == ref["t_aln"]).all())
        extra = dict(parity=ok, cpu_lean_oracle_seconds=cpu, cpu_gcups=len(q) * len(t) / cpu / 1e9)
    emit("C4 global %dx%d DNA +5/-4 d=e=8 (wavefront + device traceback, e2e single-pair call)" % (len(q), len(t)),
         float(len(q)) * len(t), secs, score=r.corner_score, aln_len=len(r.alignment.query), **extra)


def c12():
    ex = json.load(open(os.path.join(ROOT, "tests", "golden", "examples.json")))
    book = [ab.Protein.str_to_vec(r["seq"]) for r in ex["book_example_1.fasta"]]
    prot = [ab.Protein.str_to_vec(r["seq"]) for r in ex["protein.fasta"]]
    gene = [ab.DNA.from_u8_vec(r["seq"].encode()) for r in ex["human_gene_example.fasta"]]
    r, s = timed_pair("global", book[0], book[1], 8, 8, ab.get_blosum62(), ab.Protein, reps=20)
    emit("C1 book example global BLOSUM62 d=e=8 (latency of one call)", 70.0, s, score=r.corner_score)
    r, s = timed_pair("local", prot[0], prot[1], 11, 2, ab.get_blosum62(), ab.Protein, reps=10)
    emit("C2 protein.fasta local BLOSUM62 d=11 e=2 (latency of one call)", 340.0 * 341, s, f=r.alignment.f)
    r, s = timed_pair("global", gene[0], gene[1], 8, 8, ab.dna_matrix(), ab.DNA, reps=10)
    emit("C4 file human_gene_example global d=e=8 (latency of one call)", 1231.0 * 1020, s, score=r.corner_score)


def c5(P, K):
    pairs = W.protein_pairs(0, P, 270, 330)
    mats = W.candidate_matrices(K)
    freqs = W.pair_frequencies(pairs, 24)
    ab.matrix_search(pairs.__class__.from_lists(*zip(*[pairs.pair(i) for i in range(min(P, 4))])),
                     mats[:2], 5, 5, -1.0, 576.0, freqs[:min(P, 4)])  # warm-up
    runs = []
    for _ in range(3):  # the first call of a process pays lazy kernel loading and the first allocations
        t0 = time.perf_counter()
        res = ab.heuristic_batch(pairs, mats, 5, 5, -1.0, 576.0, freqs, want_matrices=False)
        runs.append(time.perf_counter() - t0)
    secs = min(runs)
    n = np.diff(pairs.q_off).astype(np.float64)
    m = np.diff(pairs.t_off).astype(np.float64)
    passes = res.passes.reshape(P, K).astype(np.float64)
    cells = float((passes * (n * m)[:, None]).sum())
    emit("C5 heuristic search %d pairs x %d matrices (~300 aa, FP64)" % (P, K), cells, secs,
         items=P * K, mean_passes=float(passes.mean()), ok_items=int((res.status == 0).sum()),
         seconds_of_each_call=runs)


def n3(length=20_000_000):
    """N3: the repeat-search engine loop (engine/calc.rs:149-241) on a synthetic chromosome with planted
    repeats: shuffle scan for mean / std, then `repeats` cycles of scan -> z filter -> align hits ->
    overlap filter -> PWM rebuild (reference defaults: 300 / 30, del 30, ext 7)."""
    from aligner_b200 import repeat_search as R
    seq, repeat, planted = W.dna_chromosome_with_repeats(length, 300, 400, seed=11, sub_rate=0.2)
    raw = np.frombuffer(b"ATCG", dtype=np.uint8)[seq].tobytes()
    opts = R.Options(repeats=3)
    start = W.pwm_from_repeat(repeat, 1.0, -1.0)
    R.perform_calculation_per_sequence(R.Options(repeats=1), raw[:200000], "warm", np.random.default_rng(0), matrix=start)
    t0 = time.perf_counter()
    res = R.perform_calculation_per_sequence(opts, raw, "chr", np.random.default_rng(1), matrix=start)
    secs = time.perf_counter() - t0
    windows = (length + 29) // 30
    cells = 4.0 * windows * 330 * 300  # one shuffled scan + three cycles
    tasks = res["direct"][0]
    emit("N3 repeat search: %d bp, 1 shuffled scan + 3 cycles of %d windows (330 x 300 FP64), hits aligned, PWM rebuilt"
         % (length, windows), cells, secs, hits=len(tasks), planted=len(planted),
         strong_hits=int(sum(t.z >= 8 for t in tasks)))


def n2():
    """N2: the p-value shuffle batch (statistics/mod.rs:255-292) and the C3 reads, score-only."""
    from aligner_b200 import statistics as S
    pp = W.protein_pairs(1, 1, 280, 320)
    query, target = pp.pair(0)
    targets = S.shuffled_targets(target, rng=np.random.default_rng(1))
    M = ab.get_blosum62()
    ab.shuffle_scores(query, target, 100.0, 11, 2, M, targets=targets[:64])  # warm-up
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        scores, lengths, _ = ab.shuffle_scores(query, target, 100.0, 11, 2, M, targets=targets)
        best = min(best, time.perf_counter() - t0)
    cells = float(len(query)) * float(sum(len(t) for t in targets))
    emit("N2 p-value batch: %d aa query x 4999 shuffled targets, BLOSUM62 d=11 e=2, score-only (e2e call)" % len(query),
         cells, best, mean_score=float(scores[1:].mean()))
    pairs = PackedPairs.from_lists([query] * len(targets), targets)
    db = ab.DeviceBatch(N.LOCAL, pairs, M, 11, 2, score_only=True)
    db.run_timed(2, profile_fill=False)
    ms, _ = db.run_timed(10, profile_fill=False)
    emit("N2 p-value batch, device-resident, score-only", cells, ms / 10 / 1e3)
    db.close()
    db = ab.DeviceBatch(N.LOCAL, pairs, M, 11, 2)
    db.run_timed(2, profile_fill=False)
    ms, _ = db.run_timed(10, profile_fill=False)
    emit("N2 p-value batch, device-resident, with traceback", cells, ms / 10 / 1e3)
    db.close()
    reads = W.dna_read_batch(0, 1000000, n=150)
    for d, e in [(10, 2), (6, 6)]:
        for so in (True, False):
            db = ab.DeviceBatch(N.LOCAL, reads, ab.dna_matrix(5, -4), d, e, score_only=so)
            db.run_timed(2, profile_fill=False)
            ms, _ = db.run_timed(5, profile_fill=False)
            emit("C3 reads 10^6 x 150x150 local d=%d e=%d device-resident, %s" % (d, e, "score-only" if so else "with traceback"),
                 1e6 * 150 * 150, ms / 5 / 1e3)
            db.close()


def n1(length=20_000_000):
    """N1: one cycle of bin/latent-repeat-search's window scan (engine/calc.rs:107-137): windows of
    repeat_length + query_offset = 330 every query_offset = 30 bases against a 4 x 300 FP64 profile."""
    from aligner_b200 import pwm as P
    import oracle_lib as O
    seq, repeat, _ = W.dna_chromosome_with_repeats(length, 300, 200, seed=9)
    pwm = W.pwm_from_repeat(repeat, 1.0, -1.0) + 0.03125 * np.random.default_rng(2).normal(0, 1, (4, 300))
    t0 = time.perf_counter()
    rs = P.ResidentSequence(seq)
    up = time.perf_counter() - t0
    rs.align_windows(pwm, 5.0, 5.0, first=0, step=30, window=330, count=1000)  # warm-up
    for d, e, tag in [(5.0, 5.0, "del == ext"), (6.0, 2.0, "stateful penalty")]:
        best = 1e9
        for _ in range(2):
            t0 = time.perf_counter()
            res = rs.align_windows(pwm, d, e, first=0, step=30, window=330)
            best = min(best, time.perf_counter() - t0)
        starts = np.arange(0, length, 30, dtype=np.uint64)
        lens = np.minimum(330, length - starts).astype(np.uint32)
        cells = float(lens.astype(np.float64).sum()) * 300
        # bounded CPU sample of the same windows on all host threads + parity of that sample
        sub = np.arange(0, len(starts), max(1, len(starts) // 4000))
        t0 = time.perf_counter()
        ref = O.pwm_windows(seq, starts[sub], lens[sub], pwm, d, e, threads=os.cpu_count() or 8, want_alignments=False)
        cpu = time.perf_counter() - t0
        emit("N1 PWM window scan: %d bp, %d windows of 330 vs 4x300 FP64 profile, %s, score-only (call incl. D2H)"
             % (length, res.count, tag), cells, best, upload_seconds=up, parity=bool((res.f[sub] == ref["f"]).all()),
             cpu_gcups=float(lens[sub].astype(np.float64).sum()) * 300 / cpu / 1e9, cpu_threads=os.cpu_count(),
             hits_z3=int((res.f > res.f.mean() + 3 * res.f.std()).sum()))
        hits = np.flatnonzero(res.f > res.f.mean() + 3 * res.f.std())
        t0 = time.perf_counter()
        al = rs.align_windows(pwm, d, e, starts=starts[hits], lens=lens[hits], want_alignments=True, want_freq=True)
        emit("N1 PWM alignments + frequency matrices of the z >= 3 hits (%d windows), %s" % (len(hits), tag),
             float(lens[hits].astype(np.float64).sum()) * 300, time.perf_counter() - t0,
             mean_aln_len=float(al.aln_len.mean()) if len(hits) else 0.0)
    rs.close()


if __name__ == "__main__":
    what = sys.argv[1:] or ["c12", "c4"]
    if N.lib().alnr_device_count() < 1:
        raise SystemExit("needs a GPU")
    if "c12" in what:
        c12()
    if "c4" in what:
        c4(20000, True)
    if "c4full" in what:
        c4(100000, "nocheck" not in what)
    if "c5" in what:
        c5(200, 50)
    if "c5big" in what:
        c5(2000, 100)
    if "n1" in what:
        n1()
    if "n3" in what:
        n3()
    if "n2" in what:
        n2()
    if "peaks" in what:
        import ctypes as C
        a, b = C.c_double(), C.c_double()
        N.lib().alnr_measure_int_peak(0, C.byref(a), C.byref(b))
        c, d = C.c_double(), C.c_double()
        N.lib().alnr_measure_fp64_peak(0, C.byref(c), C.byref(d))
        print(json.dumps(dict(int_alu_tops=a.value / 1e12, int_mixed_tops=b.value / 1e12,
                              f64_add_tops=c.value / 1e12, f64_addmax_tops=d.value / 1e12)), flush=True)
