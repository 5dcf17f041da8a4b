// wave_launch.h — host entry points of the intra-pair wavefront kernel (wave.cu).
#pragma once
#include "batch.h"

namespace alnr {

// true when a pair of m x n cells should run on the wavefront kernel (long, static penalty).
bool wave_eligible(const alnr_batch_impl &B, uint32_t m, uint32_t n);
// enqueue init + one launch per tile anti-diagonal + finish for the single pair of chunk c.
int wave_run_pair(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const FillOut &fo);

}  // namespace alnr
