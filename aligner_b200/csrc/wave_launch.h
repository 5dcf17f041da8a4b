// wave_launch.h — host entry points of the intra-pair wavefront path (wave.cu, kernels_band.cuh).
#pragma once
#include "batch.h"

namespace alnr {

// true when a pair of m x n cells should run on the wavefront kernel (long, static penalty).
bool wave_eligible(const alnr_batch_impl &B, uint32_t m, uint32_t n);
// rows per lane of the band pipeline kernel (its D layout is DLayout::make_blocked(R, ...)).
uint32_t wave_band_rows(uint32_t m);
// enqueue init + the persistent band pipeline launch + finish for the single pair of chunk c.
int wave_run_pair(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const FillOut &fo);

// checkpointed traceback (pairs whose direction matrix exceeds the scratch budget)
uint32_t wave_ckpt_group(uint32_t m, uint32_t n, uint64_t budget);
struct TraceArgs;
int wave_ckpt_trace(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const TraceArgs &ta, cudaStream_t s);

}  // namespace alnr
