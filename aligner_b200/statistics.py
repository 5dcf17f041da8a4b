"""The alignment batch of the p-value estimate.

``calculate_p_value`` (aligner-core/src/statistics/mod.rs:240-307) draws SEQUENCES-1 = 4999
shuffled, randomly shortened copies of the target (``shuffle_and_randomize_sequence``,
:309-320), aligns the query to each with ``SimpleLocalAligner`` on 10 OS threads (:255-286) and
keeps only ``alignment.f`` and the copy's length; the (K, lambda, H) fit that follows
(:36-123) is scalar host math and stays outside this package.  Here the 4999 alignments are
ONE score-only batch on the GPU (``alnr_score_batch``: no direction scratch, no traceback).

The reference shuffles with ``thread_rng`` (not reproducible), so parity is defined on the scores
of a GIVEN set of shuffled targets: tests feed the same targets to the oracle.
"""
import numpy as np

from . import _native as N
from .batch import PackedPairs, score_batch

THREADS = 10        # statistics/mod.rs:10
SEQUENCES = 5000    # statistics/mod.rs:11


def shuffle_count(sequences=SEQUENCES, threads=THREADS):
    """Number of shuffles the reference runs: every thread does SEQUENCES/THREADS, except thread 5
    which does SEQUENCES - SEQUENCES/THREADS*(THREADS-1) - 1 (statistics/mod.rs:263-266)."""
    per = sequences // threads
    return per * (threads - 1) + (sequences - per * (threads - 1) - 1)


def shuffled_targets(target, count=None, rng=None):
    """``count`` draws of shuffle_and_randomize_sequence (statistics/mod.rs:309-320): drop the last
    ``lock`` symbols, lock uniform in 0..7, then shuffle what is left."""
    rng = np.random.default_rng() if rng is None else rng
    target = np.asarray(target, dtype=np.uint8)
    count = shuffle_count() if count is None else count
    out = []
    for _ in range(count):
        lock = int(rng.integers(0, 7))
        out.append(rng.permutation(target[:len(target) - lock]))
    return out


def shuffle_scores(query, target, initial_score, del_, ins, matrix, targets=None, rng=None,
                   device=-1, numeric=N.NUM_AUTO):
    """The inputs of ``calculate_distribution_params`` as the reference builds them
    (statistics/mod.rs:250-251, :288-292): scores = [initial_score, f of every shuffle],
    lengths = [len(target), len of every shuffle].  ``targets`` overrides the random draws.
    Returns (scores float64[S], lengths uint64[S], ScoreResult of the batch)."""
    query = np.asarray(query, dtype=np.uint8)
    target = np.asarray(target, dtype=np.uint8)
    if targets is None:
        targets = shuffled_targets(target, rng=rng)
    pairs = PackedPairs.from_lists([query] * len(targets), targets)
    res = score_batch(N.LOCAL, pairs, matrix, del_, ins, device=device, numeric=numeric)
    bad = np.flatnonzero(res.status != 0)  # the reference unwraps every alignment (:268-274)
    if len(bad):
        raise N.AlignerError(int(res.status[bad[0]]), "shuffle %d" % int(bad[0]))
    scores = np.concatenate([[float(initial_score)], res.f])
    lengths = np.concatenate([[len(target)], [len(t) for t in targets]]).astype(np.uint64)
    return scores, lengths, res
