// kernels_strip.cu — s16x2 strip kernel family (placeholder: not yet eligible for any batch).
#include "strip_launch.h"

namespace alnr {

bool strip_eligible(const alnr_batch_impl &, double) { return false; }
uint32_t d_rows_per_word(const alnr_batch_impl &, uint32_t) { return 16; }
int strip_prepare(alnr_batch_impl &) { return ALNR_OK; }
int strip_run_chunk(alnr_batch_impl &, const Chunk &, const ChunkView &, const DLayout &,
                    const FillOut &) {
  set_error_detail("strip kernel not built");
  return ALNR_CALCULATION_ERROR;
}

}  // namespace alnr
