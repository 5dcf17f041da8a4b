// strip_launch.h — host entry points of the s16x2 strip kernel family (kernels_strip.cu).
#pragma once
#include "batch.h"

namespace alnr {

// true when the batch can run in packed 16-bit lanes (two pairs per thread) exactly.
bool strip_eligible(const alnr_batch_impl &B, double value_bound);
// rows per 32-bit D word for a chunk whose longest target has m rows (<= 16).
uint32_t d_rows_per_word(const alnr_batch_impl &B, uint32_t m);
// pairs per resident wave of the strip kernel (chunk sizes are multiples of it).
uint32_t strip_wave_pairs();
// words of one boundary-row buffer (per chunk parity the batch keeps two of them, see kernels_strip.cu)
uint64_t strip_row_words(const alnr_batch_impl &B);
// one-time per-batch preparation (packed sequences, profile table).
int strip_prepare(alnr_batch_impl &B);
// phase 1 (B.stream): pack the chunk's sequences and run the speculative main pass.
int strip_main_pass(alnr_batch_impl &B, const Chunk &c, const ChunkView &cv, const DLayout &dl,
                    const FillOut &fo);
// phase 2 (B.s2; local mode with del != ext only): re-run passes over the fail list and the exact
// column-major kernel for pairs that never settle.
int strip_rerun_passes(alnr_batch_impl &B, const Chunk &c, const ChunkView &cv, const DLayout &dl,
                       const FillOut &fo);

// diagnostic: pairs on the re-run list after pass 0..3, summed over chunks (synchronises).
int strip_fail_counts(alnr_batch_impl &B, uint64_t out[4]);

}  // namespace alnr
