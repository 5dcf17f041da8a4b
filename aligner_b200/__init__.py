"""aligner_b200 — B200-native replacement for aligner-core's pairwise alignment hot path.

Host-side mirror of the reference API (``SimpleGlobalAligner``, ``SimpleLocalAligner``,
``Heuristics``, ``Alignment``, ``AlignmentResult``, ``Protein``, ``DNA``, ``get_blosum62``) over the
C-ABI CUDA library ``libaligner_b200.so`` (include/aligner_b200.h).  No CPU fallback exists.
"""
from . import _native
from ._native import AlignerError, GLOBAL, LOCAL, NUM_AUTO, NUM_F64, NUM_I16X2, NUM_I32
from .aligners import (Alignment, AlignmentResult, Heuristics, SimpleGlobalAligner,
                       SimpleLocalAligner)
from .batch import (BatchResult, CompactResult, DeviceBatch, Packed2Pairs, PackedPairs, ScoreResult, align_batch,
                    pack2bit, score_batch)
from .enums import DNA, Direction, Error, Protein
from .heuristic import HeuristicAligner, heuristic_batch, matrix_search, transform_matrix
from .matrices import dna_matrix, get_blosum62
from .pwm import (HeuristicPWMAligner, PWMAligner, PWMAlignment, ResidentSequence,
                  get_random_pwm)
from .statistics import shuffle_scores, shuffled_targets
from .sharding import (align_batch_sharded, matrix_search_sharded, pwm_scan_sharded, score_batch_sharded,
                       shard_range)

__all__ = [
    "AlignerError", "Alignment", "AlignmentResult", "BatchResult", "CompactResult", "Packed2Pairs", "pack2bit", "DNA", "DeviceBatch", "Direction",
    "Error", "GLOBAL", "HeuristicAligner", "Heuristics", "heuristic_batch", "matrix_search", "transform_matrix", "LOCAL", "NUM_AUTO", "NUM_F64", "NUM_I16X2", "NUM_I32",
    "PackedPairs", "Protein", "SimpleGlobalAligner", "SimpleLocalAligner", "align_batch",
    "dna_matrix", "get_blosum62", "ScoreResult", "score_batch", "PWMAligner", "PWMAlignment", "HeuristicPWMAligner", "ResidentSequence", "get_random_pwm", "shuffled_targets", "shuffle_scores", "align_batch_sharded", "matrix_search_sharded", "shard_range", "pwm_scan_sharded", "score_batch_sharded",
]
