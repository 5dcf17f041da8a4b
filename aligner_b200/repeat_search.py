"""Dispersed-repeat search on top of the PWM window scan.

Host-side mirror of ``aligner-core/src/bin/latent-repeat-search``: the engine loop
(``engine/calc.rs:19-241``), the overlap filter and coordinate helpers (``engine/mod.rs:49-152``),
the DNA import with skipped-symbol indices (``aligner-core/src/enums.rs:489-527``), the exploring /
csv / testing commands (``cmd/*.rs``) and the CSV + JSON outputs (``main.rs:42-64``,
``aligner-helpers/src/csv/mod.rs:8-56``).  Every alignment runs on the GPU: a cycle is ONE strided
score-only scan of the HBM-resident chromosome plus ONE call that aligns the z >= 3 windows
(``ResidentSequence.align_windows``); nothing here computes a DP cell.

The reference draws its random PWM, the shuffle and the mutations from ``thread_rng`` and gathers
results from worker threads in arrival order; here every random draw comes from the ``rng``
argument and sums run in window order, so a run is reproducible.
"""
import json
import math

import numpy as np

from . import _native as N
from .enums import DNA, Error
from .heuristic import transform_matrix
from .pwm import HeuristicPWMAligner, PWMAlignment, ResidentSequence, get_random_pwm
from .aligners import Heuristics

Z = 3.0  # engine/calc.rs:17
DIRECT, INVERSE = "direct", "inverse"  # cmd/mod.rs:15-16


class Index:
    """aligner-core/src/enums.rs:567-572"""

    def __init__(self, coord, offset, local_offset):
        self.coord, self.offset, self.local_offset = coord, offset, local_offset

    def __eq__(self, o):
        return (self.coord, self.offset, self.local_offset) == (o.coord, o.offset, o.local_offset)

    def __repr__(self):
        return "Index(coord=%d, offset=%d, local_offset=%d)" % (self.coord, self.offset, self.local_offset)


class Task:
    """engine/task.rs:4-16; equality is by left_coord only (:12-16)."""

    def __init__(self, alignment, left_coord, right_coord, z):
        self.alignment, self.left_coord, self.right_coord, self.z = alignment, left_coord, right_coord, z

    def __eq__(self, o):
        return self.left_coord == o.left_coord

    def __repr__(self):
        return "Task(z=%r, left_coord=%d, right_coord=%d)" % (self.z, self.left_coord, self.right_coord)


class Options:
    """cmd/mod.rs:20-48 with the defaults of args.rs:3-44."""

    def __init__(self, repeat_length=300, query_offset=30, deletions=30.0, extension=7.0, rsquared=100000.0,
                 kd=0.0, threads=1, repeats=10, simple_init=False, reverse=False):
        self.repeat_length, self.query_offset = int(repeat_length), int(query_offset)
        self.deletions, self.extension = float(deletions), float(extension)
        self.rsquared, self.kd = float(rsquared), float(kd)
        self.threads, self.repeats = int(threads), int(repeats)
        self.simple_init, self.reverse = bool(simple_init), bool(reverse)


def dna_from_u8_with_freqs_and_indices(raw):
    """DNA::from_u8_vec_with_freqs_and_indices (enums.rs:489-527): bytes other than A/T/C/G are
    skipped; every run of skipped bytes leaves an Index so that window coordinates can be mapped back
    to the raw sequence.  Returns (codes uint8, freqs float64[4], indices newest first)."""
    raw = np.frombuffer(bytes(raw), dtype=np.uint8)
    lut = np.full(256, 255, dtype=np.uint8)
    for ch, code in zip(b"ATCG", range(4)):
        lut[ch] = code
    if ((raw == ord("_")) | (raw == ord("+"))).any():
        raise Error("ValidationError", "'_' / '+' in the input (the reference indexes freqs out of bounds)")
    codes_all = lut[raw]
    ok = codes_all != 255
    codes = codes_all[ok]
    freqs = np.bincount(codes, minlength=4).astype(np.float64)
    indices = []
    bad = ~ok
    # a run of skipped bytes ending right before raw position i (a kept byte) -> Index at i
    starts_kept_after_bad = np.flatnonzero(ok[1:] & bad[:-1]) + 1
    cum_bad = np.cumsum(bad)
    prev = 0
    for i in starts_kept_after_bad:
        count = int(cum_bad[i])
        indices.append(Index(int(i) - count, count, count - prev))
        prev = count
    indices.reverse()                                   # enums.rs:522
    freqs = freqs / float(len(codes)) if len(codes) else freqs * float("nan")  # :524
    return codes, freqs, indices


def index_coord(target, indices):
    """engine/mod.rs:121-129"""
    for index in indices:
        if target >= index.coord:
            return target + index.offset
    return target


def rotate_indices(indices, query_length):
    """engine/mod.rs:131-152"""
    if not indices:
        return []
    ref = indices[0]
    full_length = query_length + ref.offset
    offset = 0
    result = []
    for index in indices:
        offset += index.local_offset
        result.append(Index(full_length - index.coord - ref.offset, offset, index.local_offset))
    result.reverse()
    return result


def check_intersection(c1, c2):
    """engine/mod.rs:102-119 (the fourth test repeats the first, as in the reference)"""
    if c2[0] <= c1[0] <= c2[1]:
        return True
    if c2[0] <= c1[1] <= c2[1]:
        return True
    if c2[0] >= c1[0] and c2[1] <= c1[1]:
        return True
    return False


def filter_tasks(tasks):
    """engine/mod.rs:49-100: sort by left_coord, then repeatedly take the first task and the run of
    following tasks that intersect IT, keep the one with the largest z (the last of equal maxima, as
    Iterator::max_by does).  When the run reaches the end of the list the last task is looked at
    again (index + 1 points at it) and is pushed unless an equal left_coord is already kept."""
    tasks = list(tasks)
    if len(tasks) <= 1:
        return tasks
    result = []
    tasks.sort(key=lambda t: t.left_coord)  # stable, like slice::sort_by
    while tasks:
        if len(tasks) == 1:
            if tasks[0] not in result:
                result.append(tasks[0])
            break
        current = tasks[0]
        batch = [current]
        index = 0
        for i, task in enumerate(tasks[1:]):
            index = i
            if check_intersection((current.left_coord, current.right_coord), (task.left_coord, task.right_coord)):
                batch.append(task)
            else:
                break
        if len(batch) == 1:
            result.append(batch[0])
        else:
            best = batch[0]
            for t in batch[1:]:
                if not (t.z < best.z):  # max_by returns the last maximum
                    best = t
            result.append(best)
        tasks = tasks[index + 1:]
    return result


def _window_starts(length, first_mult, step, opts):
    """The j values of the worker loops (calc.rs:50, :110): thread i walks
    (i*query_offset .. length).step_by(step * threads); listed thread by thread."""
    if step * opts.threads <= 0:
        raise ValueError("step_by(0): sequence shorter than 1000 symbols with simple_init")
    out = []
    for i in range(opts.threads):
        out.append(np.arange(i * first_mult, length, step * opts.threads, dtype=np.uint64))
    return np.concatenate(out) if out else np.zeros(0, dtype=np.uint64)


def _borders(starts, length, opts):
    """border = j + repeat_length + query_offset, or length when that reaches it (calc.rs:51-55)"""
    w = opts.repeat_length + opts.query_offset
    s = starts.astype(np.int64)
    return np.where(s + w >= length, length, s + w).astype(np.int64)


class GpuBackend:
    """The PWM aligner calls of the engine, on one HBM-resident sequence."""

    def __init__(self, codes, device=-1):
        self.seq = ResidentSequence(codes, device)

    def scores(self, pwm, del_, ext, starts, lens):
        return self.seq.align_windows(pwm, del_, ext, starts=starts, lens=lens).f

    def alignments(self, pwm, del_, ext, starts, lens):
        res = self.seq.align_windows(pwm, del_, ext, starts=starts, lens=lens, want_alignments=True)
        return [res.alignment(i) for i in range(res.count)]

    def close(self):
        self.seq.close()


def calculate_starting_values(query, matrix, opts, rng, backend=GpuBackend):
    """calc.rs:19-86: mean and (population) standard deviation of f over windows of a SHUFFLED copy."""
    shuffled = rng.permutation(np.asarray(query, dtype=np.uint8))
    length = len(shuffled)
    step = length // 1000 if opts.simple_init else opts.query_offset
    starts = _window_starts(length, opts.query_offset, step, opts)
    lens = (_borders(starts, length, opts) - starts.astype(np.int64)).astype(np.uint32)
    be = backend(shuffled)
    try:
        fs = be.scores(matrix, opts.deletions, opts.extension, starts, lens)
    finally:
        be.close()
    mean = _seq_sum(fs) / len(fs)
    std = math.sqrt(_seq_sum((fs - mean) * (fs - mean)) / len(fs))
    return mean, std


def _seq_sum(values):
    """Iterator::sum::<f64>(): a left-to-right fold."""
    v = np.asarray(values, dtype=np.float64)
    return float(np.cumsum(v)[-1]) if len(v) else 0.0  # cumsum is a sequential fold


def calculate_cycle(be, length, matrix, indices, mean, std, opts):
    """calc.rs:88-147 on an open backend: every window, z = (f - mean) / std, tasks with z >= 3."""
    starts = _window_starts(length, opts.query_offset, opts.query_offset, opts)
    borders = _borders(starts, length, opts)
    lens = (borders - starts.astype(np.int64)).astype(np.uint32)
    f = be.scores(matrix, opts.deletions, opts.extension, starts, lens)
    with np.errstate(divide="ignore", invalid="ignore"):  # std == 0 gives inf / NaN as in the reference
        z = (f - mean) / std
    hits = np.flatnonzero(z >= Z)
    if len(hits) == 0:
        return []
    alns = be.alignments(matrix, opts.deletions, opts.extension, starts[hits], lens[hits])
    return [Task(a, index_coord(int(starts[h]), indices), index_coord(int(borders[h]), indices), float(z[h]))
            for a, h in zip(alns, hits)]


def perform_calculation_per_sequence(opts, raw_seq, head="", rng=None, backend=GpuBackend, matrix=None):
    """calc.rs:149-241.  Returns {"direct": (tasks, matrix)[, "inverse": (tasks, matrix)]}."""
    rng = np.random.default_rng() if rng is None else rng
    query, frequencies, indices = dna_from_u8_with_freqs_and_indices(raw_seq)
    if matrix is None:
        matrix = get_random_pwm(opts.repeat_length, rng)
    r2 = opts.deletions * opts.extension
    matrix = transform_matrix(matrix, 0.0, r2, frequencies)                 # :157-163
    mean, std = calculate_starting_values(query, matrix, opts, rng, backend)  # :173
    result = {}
    tasks = []
    be = backend(query)
    try:
        for i in range(opts.repeats):
            new_tasks = calculate_cycle(be, len(query), matrix, indices, mean, std, opts)
            if not new_tasks:
                break
            tasks = filter_tasks(new_tasks)
            if i < opts.repeats - 1:
                fs = [t.alignment.f for t in tasks]
                mean = _seq_sum(fs) / len(tasks)
                std = _seq_sum([(v - mean) * (v - mean) for v in fs]) / len(tasks)  # :199-203: no sqrt
                m = np.zeros_like(matrix)
                for t in tasks:
                    m = m + t.alignment.get_frequency_matrix()
                matrix = transform_matrix(m, 0.0, r2, frequencies)          # :210-216
    finally:
        be.close()
    result[DIRECT] = (tasks, matrix.copy())
    if opts.reverse:                                                        # :223-238
        rq = query[::-1].copy()
        rotated = rotate_indices(indices, len(rq))
        be = backend(rq)
        try:
            inv = calculate_cycle(be, len(rq), matrix, rotated, mean, std, opts)
        finally:
            be.close()
        result[INVERSE] = (filter_tasks(inv), matrix)
    return result


# ---- commands (cmd/exploring.rs, cmd/csv.rs, cmd/testing.rs) and outputs (main.rs) ---------------
def import_fasta(path):
    """engine/sequences.rs:10-32: [(head, raw sequence bytes)]; an empty file is an error."""
    records, head, chunks = [], None, []
    with open(path, "rb") as fh:
        for line in fh:
            line = line.rstrip(b"\r\n")
            if line.startswith(b">"):
                if head is not None:
                    records.append((head, b"".join(chunks)))
                head, chunks = line[1:].decode(), []
            elif head is not None:
                chunks.append(line)
    if head is not None:
        records.append((head, b"".join(chunks)))
    if not records:
        raise ValueError("empty fasta file")
    return records


def read_csv(path):
    """aligner-helpers/src/csv/mod.rs:16-39: {name: [(z_value, left_coord, right_coord)]}"""
    import csv
    out = {}
    with open(path, newline="") as fh:
        for row in csv.DictReader(fh):
            out.setdefault(row["name"], []).append((float(row["z_value"]), int(row["left_coord"]), int(row["right_coord"])))
    return out


def prepare_sequence(raw_seq, records):
    """engine/sequences.rs:34-43: found repeats are masked with 'N' (skipped by the DNA import)."""
    buf = bytearray(raw_seq)
    for _z, left, right in records:
        buf[left:right] = b"N" * (right - left)
    return bytes(buf)


def run_exploring(opts, fasta_path, csv_path=None, rng=None, backend=GpuBackend):
    """cmd/exploring.rs:6-33 and, with ``csv_path``, cmd/csv.rs:8-49."""
    rng = np.random.default_rng() if rng is None else rng
    masks = read_csv(csv_path) if csv_path else {}
    result = {}
    for head, raw in import_fasta(fasta_path):
        if head in masks:
            raw = prepare_sequence(raw, masks[head])
        r = perform_calculation_per_sequence(opts, raw, head, rng, backend)
        if DIRECT in r:
            result[head] = r[DIRECT]
        if INVERSE in r:
            result["%s-reversed" % head] = r[INVERSE]
    return result


def mutate(sequence, offset, start, rng):
    """engine/mod.rs:33-47: every offset-th symbol from `start` is redrawn uniformly."""
    out = np.array(sequence, dtype=np.uint8)
    idx = np.arange(start, len(out), offset)
    out[idx] = rng.integers(0, 4, len(idx))
    return out


def run_testing(opts, rng=None, backend=GpuBackend, sequence_length=100000, descendants=10):
    """cmd/testing.rs:13-75: plant mutated copies of a random sample in a random sequence, fit a PWM
    to the sample with HeuristicPWMAligner, run one cycle."""
    rng = np.random.default_rng() if rng is None else rng
    raw = rng.integers(0, 4, sequence_length).astype(np.uint8)
    query = rng.integers(0, 4, opts.repeat_length + opts.query_offset).astype(np.uint8)
    freqs = np.bincount(query, minlength=4) / float(len(query))
    res = HeuristicPWMAligner.from_seqs(query).perform_alignment(
        opts.deletions, opts.extension, get_random_pwm(opts.repeat_length, rng),
        Heuristics(opts.kd, opts.rsquared, freqs), want_matrices=False)
    matrix = res.matrix
    kids = [mutate(query, 4, i, rng) for i in range(descendants)]  # MutationPercent::Quarter
    off = len(raw) // (len(kids) + 1)
    parts = [raw[:off]]
    for i, kid in enumerate(kids):  # :50-54 (the slice re-uses offset*i .. offset*(i+1))
        parts += [kid, raw[off * i:off * (i + 1)]]
    sequence = np.concatenate(parts)
    mean, std = calculate_starting_values(sequence, matrix, opts, rng, backend)
    be = backend(sequence)
    try:
        tasks = calculate_cycle(be, len(sequence), matrix, [], mean, std, opts)
    finally:
        be.close()
    return {"test": (tasks, matrix)}


def write_outputs(result, output_path="output.csv", matrices_path=None):
    """main.rs:42-64: one CSV row per task (Record: name, z_value, left_coord, right_coord) and the
    matrices as JSON in ndarray's serde layout {"v":1,"dim":[rows,cols],"data":[...]}."""
    import csv
    matrices_path = matrices_path or "%s.matrices.json" % output_path
    with open(output_path, "w", newline="") as fh:
        w = csv.writer(fh, lineterminator="\n")
        w.writerow(["name", "z_value", "left_coord", "right_coord"])
        for key, (tasks, _m) in result.items():
            for t in tasks:
                w.writerow([key, repr(float(t.z)), t.left_coord, t.right_coord])
    mats = {k: {"v": 1, "dim": list(m.shape), "data": [float(x) for x in np.asarray(m).reshape(-1)]}
            for k, (_t, m) in result.items()}
    with open(matrices_path, "w") as fh:
        json.dump(mats, fh, separators=(",", ":"))
    return output_path, matrices_path


def main(argv=None):
    """args.rs:3-44 / main.rs:18-73"""
    import argparse
    ap = argparse.ArgumentParser(prog="latent-repeat-search (aligner_b200)")
    ap.add_argument("-i", "--input")
    ap.add_argument("-o", "--output")
    ap.add_argument("--csv")
    ap.add_argument("-d", "--deletions", type=float, default=30.0)
    ap.add_argument("-e", "--extension", type=float, default=7.0)
    ap.add_argument("--rsquared", type=float, default=100000.0)
    ap.add_argument("--kd", type=float, default=0.0)
    ap.add_argument("-q", "--query-offset", type=int, default=30)
    ap.add_argument("-r", "--repeat-length", type=int, default=300)
    ap.add_argument("--threads", type=int, default=1)
    ap.add_argument("--simple-init", action="store_true")
    ap.add_argument("--repeats", type=int, default=10)
    ap.add_argument("--reverse", action="store_true")
    ap.add_argument("--seed", type=int, default=None, help="seed of every random draw (not in the reference)")
    a = ap.parse_args(argv)
    opts = Options(a.repeat_length, a.query_offset, a.deletions, a.extension, a.rsquared, a.kd, a.threads,
                   a.repeats, a.simple_init, a.reverse)
    rng = np.random.default_rng(a.seed)
    result = run_testing(opts, rng) if a.input is None else run_exploring(opts, a.input, a.csv, rng)
    out = a.output or "output.csv"
    mats = "%s.matrices.json" % a.output if a.output else "matrices.json"
    paths = write_outputs(result, out, mats)
    print("\nOutput written to:\n 1. Result: %s\n 2. Matrices: %s" % paths)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
