// kernels_band.cuh — argument block and result kernel shared by the long-pair (intra-pair wavefront)
// path.  The first-generation kernel that lived here (one 4 x 1 column per lane and step, 8 % of the
// INT32 roofline) was replaced by the 4 x 4 tile pipeline of kernels_tile.cuh and removed in round 2.
//
// Static penalty: global mode (cell (1,1) pays del, every other cell ext — get_direction never
// returns Beginning, simple/mod.rs:88-92) and local mode with del == ext.
// Local mode with del != ext (simple/mod.rs:201-205): a cell pays del iff the cell BEFORE it in the
// reference's column-major order was a Beginning, i.e. had H == 0.  For rows y > 1 that is the
// cell above — a value the wavefront holds anyway, so the rule costs one compare + select per cell
// and is exact.  Row 1 looks back at the LAST cell of the previous column, (m, x-1), which the
// wavefront reaches ~m steps later: the pipeline speculates (delmap: the columns of row 1 that pay
// del, initially column 1 only), records the zeros of row m (zeromap), and the pass is repeated
// with the corrected map until the map reproduces itself (wave.cu; passes after convergence
// return at once, nothing synchronises with the host).
#pragma once
#include "common.cuh"

namespace alnr {


struct BandArgs {
  const uint8_t *q, *t;   // this pair's symbol codes
  uint32_t n, m;
  int V;
  const int *matrix;      // V*V ints, [t][q]
  int del, ext;
  uint32_t nbands;
  int *rows;              // [nbands][n + 1]: bottom row of every band (index x)
  uint32_t *prog;         // [nbands]: columns of the band's bottom row that are published
  uint32_t *D;            // blocked layout
  uint32_t dcols;         // words per 4-row strip
  unsigned long long *best;  // local: packed (value | ~y | ~x) maximum
  int *corner_out;        // H[m][n]
  // Checkpointed traceback (tile kernel only; wave.cu): a RE-FILL computes the bands
  // [band0, band_end) again, reading every band's top boundary row from the rows the first pass
  // kept (all final: no waiting, the bands run side by side) and writing their directions to a
  // scratch that holds only this group's strips (strip index - strip0).  ckpt != nullptr: the launch
  // first looks at the walk's state and returns when the path does not cross the group.
  uint32_t band0 = 0, band_end = 0;   // band_end == 0: all bands (first pass)
  uint32_t strip0 = 0;
  const uint32_t *ckpt = nullptr;     // CkptState words: [0] cy, [1] cx, [4] done
};
// Stateful local pairs (tile kernel, STATEFUL instantiation) reuse `prog` as their state block —
// BandArgs itself must not grow: the size of the parameter block changes ptxas' schedule of every
// instantiation (measured: three more pointers, 16.0 -> 17.6 ms on the 10^5 x 10^5 global fill).
//   prog[0] converged, prog[1] passes run, prog[4 ..] delmap, then zeromap, stateful_map_words(n) each;
//   bit c of a map = 0-based column c: delmap — cell (1, c+1) pays del; zeromap — H[m][c+1] == 0.
__host__ __device__ __forceinline__ uint32_t stateful_map_words(uint32_t n) { return (n + 31) / 32 + 4; }
__device__ __forceinline__ uint32_t *stateful_delmap(const BandArgs &a) { return a.prog + 4; }
__device__ __forceinline__ uint32_t *stateful_zeromap(const BandArgs &a) { return a.prog + 4 + stateful_map_words(a.n); }

__device__ __forceinline__ unsigned long long band_key(int v, uint32_t y, uint32_t x) {
  // larger value first, then smaller y, then smaller x (row-major first maximum, simple/mod.rs:212)
  return ((unsigned long long)(uint32_t)v << 40) | ((unsigned long long)(0xFFFFFu - y) << 20) |
         (unsigned long long)(0xFFFFFu - x);
}

// Writes the per-pair results of a long-pair fill (corner score, arg-max) into the batch's arrays.
static __global__ void band_finish_kernel(BandArgs a, int local, FillOut fo, uint32_t pair) {
  fo.corner[pair] = (double)*a.corner_out;
  if (!local) {
    fo.f[pair] = 0.0;
    fo.start[2 * (size_t)pair] = a.m;
    fo.start[2 * (size_t)pair + 1] = a.n;
  } else {
    const unsigned long long key = *a.best;
    const uint32_t y = key ? 0xFFFFFu - (uint32_t)((key >> 20) & 0xFFFFFu) : 0;
    const uint32_t x = key ? 0xFFFFFu - (uint32_t)(key & 0xFFFFFu) : 0;
    fo.f[pair] = (double)(uint32_t)(key >> 40);
    fo.start[2 * (size_t)pair] = y;
    fo.start[2 * (size_t)pair + 1] = x;
    if (!key) fo.status[pair] = 11;  // ALNR_NO_POSITIVE_CELL
    // the speculated penalty map of row 1 did not reproduce itself within the pass budget
    if (local == 2 && !a.prog[0]) fo.status[pair] = 6;  // ALNR_CALCULATION_ERROR
  }
}

}  // namespace alnr
