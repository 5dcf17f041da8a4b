// kernels_wave.cuh — intra-pair anti-diagonal wavefront fill for ONE long pair (BASELINE
// config 4: ~10^5 x 10^5 global alignment, 10^10 cells).
//
// The matrix is cut into tiles of TH = 32*R rows by TW columns.  One WARP owns a tile: lane l
// keeps R consecutive rows in registers and runs l columns behind lane l-1, so the bottom cell of
// lane l-1 arrives by one __shfl_up per column (skewed systolic sweep; 31 bubble steps per
// tile).  Tiles on the same anti-diagonal are independent: one launch per tile diagonal, no
// inter-CTA waiting.  Tile boundaries travel through three small L2-resident buffers that are
// updated in place: rowbuf[x] = H[bottom of the tile above][x], colbuf[y] = H[y][right edge of
// the tile to the left], corner[i] = H[top edge of tile row i][last finished column].
// Directions go to the 2-bit D scratch (contiguous single-pair layout, 16 rows per word; a lane
// stores its R rows as one byte (R=4) or half-word (R=8)).
//
// Valid for a STATIC penalty only: global mode (cell (1,1) pays del, every other cell ext —
// get_direction never returns Beginning, simple/mod.rs:88-92) and local mode with del == ext.
// Local mode with del != ext carries the penalty from cell (m, x-1) to (1, x)
// (simple/mod.rs:201-205), which serialises columns: such long pairs use the exact kernel.
#pragma once
#include "common.cuh"

namespace alnr {

struct WaveArgs {
  const uint8_t *q, *t;   // this pair's symbol codes
  uint32_t n, m;
  int V;
  const int *matrix;      // V*V ints, [t][q]
  int del, ext;
  int local;              // 0 global, 1 local (del == ext)
  uint32_t TW;            // tile width (multiple of 32)
  uint32_t tile_rows, tile_cols;
  uint32_t diag;          // tile anti-diagonal of this launch
  int *rowbuf;            // [n+1]
  int *colbuf;            // [m+1]
  int *corner;            // [tile_rows]
  uint8_t *Dbytes;        // D scratch as bytes
  uint64_t ncols;         // words per strip (DLayout.ncols)
  unsigned long long *best;  // local: packed (value | ~y | ~x) maximum
  int *corner_out;        // H[m][n]
};

// H[y][0], H[0][x], corners: simple/mod.rs:59-70 (global), zeros (local)
static __global__ void wave_init_kernel(WaveArgs a, uint32_t TH) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= a.n) {
    int v = 0;
    if (!a.local && i > 0) v = (i == a.n) ? -(int)(a.n + 1) * a.del : -(int)i * a.del;
    a.rowbuf[i] = v;
  }
  if (i <= a.m) {
    int v = 0;
    if (!a.local && i > 0) v = (i == a.m) ? -(int)(a.m + 1) * a.del : -(int)i * a.del;
    a.colbuf[i] = v;
  }
  if (i < a.tile_rows) {
    const uint32_t y0 = i * TH;  // H[y0][0]; the row-m quirk never sits on a tile's top edge < m
    a.corner[i] = (a.local || y0 == 0) ? 0 : -(int)y0 * a.del;
  }
  if (i == 0) { *a.best = 0ull; *a.corner_out = 0; }
}

__device__ __forceinline__ unsigned long long wave_key(int v, uint32_t y, uint32_t x) {
  // larger value first, then smaller y, then smaller x (row-major first maximum)
  return ((unsigned long long)(uint32_t)v << 40) | ((unsigned long long)(0xFFFFFu - y) << 20) |
         (unsigned long long)(0xFFFFFu - x);
}

template <int R, bool LOCAL>
__global__ void __launch_bounds__(128) fill_wave_kernel(WaveArgs a) {
  extern __shared__ int wsm[];
  int *smat = wsm;                                    // V*V
  uint8_t *sq_all = reinterpret_cast<uint8_t *>(wsm + a.V * a.V);
  for (int i = threadIdx.x; i < a.V * a.V; i += blockDim.x) smat[i] = a.matrix[i];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t *sq = sq_all + warp * a.TW;
  constexpr uint32_t TH = 32 * R;
  const uint32_t n = a.n, m = a.m;

  // tiles (i, j) with i + j == diag
  const uint32_t i_lo = a.diag >= a.tile_cols ? a.diag - a.tile_cols + 1 : 0;
  const uint32_t i_hi = min(a.diag, a.tile_rows - 1);
  const uint32_t ti = i_lo + blockIdx.x * 4 + warp;
  const bool have = ti <= i_hi;
  const uint32_t tj = have ? a.diag - ti : 0;
  const uint32_t y0 = ti * TH, x0 = tj * a.TW;
  const uint32_t TWv = have ? min(a.TW, n - x0) : 0;
  if (have) for (uint32_t k = lane; k < TWv; k += 32) sq[k] = a.q[x0 + k];
  __syncthreads();
  if (!have) return;

  const uint32_t yl0 = y0 + lane * R;  // rows yl0+1 .. yl0+R (1-based)
  const int *srow[R];                  // smat row of each target symbol
  int h[R];
  bool rvalid[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const uint32_t y = yl0 + r + 1;
    rvalid[r] = y <= m;
    srow[r] = smat + (rvalid[r] ? (int)a.t[y - 1] : 0) * a.V;
    h[r] = rvalid[r] ? a.colbuf[y] : 0;  // H[y][x0]
  }
  int diag0 = (lane == 0) ? a.corner[ti] : a.colbuf[min(yl0, m)];  // H[yl0][x0]
  int bottom = 0, rowchunk = 0;
  int best_v = 0;
  uint32_t best_y = 0, best_x = 0;
  const uint32_t steps = TWv + 31;
  const int ext = a.ext;
  // the one cell that pays `del` in global mode: (1, 1)  (simple/mod.rs:72,88-92)
  const bool first_cell_here = !LOCAL && ti == 0 && tj == 0 && lane == 0;
  // byte address of this lane's rows inside the D word of (strip, column): 16 rows per word
  uint8_t *dp = a.Dbytes + ((uint64_t)(yl0 >> 4) * a.ncols + x0) * 4 + ((yl0 & 15) >> 2) - 4 * (int64_t)lane;
  int *rowout = a.rowbuf + x0 + 1 - (int)lane;
  const bool store_d = yl0 < m;
  int up_last = 0;

  for (uint32_t s = 0; s < steps; ++s, dp += 4) {
    if ((s & 31) == 0) {  // next 32 top-boundary values, one coalesced load
      const uint32_t xx = x0 + s + lane + 1;
      rowchunk = xx <= n ? a.rowbuf[xx] : 0;
    }
    int up = __shfl_up_sync(0xffffffffu, bottom, 1);
    const int top0 = __shfl_sync(0xffffffffu, rowchunk, s & 31);
    if (lane == 0) up = top0;
    const uint32_t c = s - lane;  // wraps for s < lane
    if (c < TWv) {
      const int qs = sq[c];
      int diag = diag0;
      diag0 = up;
      up_last = up;
      int hup = up;
      uint32_t bits = 0;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int pen = (r == 0 && first_cell_here && c == 0) ? a.del : ext;
        const int sc = srow[r][qs];
        const int top = hup - pen, left = h[r] - pen, dg = diag + sc;
        const int mx = max(max(top, left), dg);
        uint32_t code = (mx == left) ? DIR_LEFT : DIR_DIAGONAL;  // enums.rs:23/:41
        code = (mx == top) ? DIR_TOP : code;                     // enums.rs:21/:39
        if (LOCAL) code = (mx == 0) ? DIR_BEGINNING : code;      // enums.rs:37
        bits |= code << (2 * r);
        diag = h[r];
        h[r] = mx;
        hup = mx;
        if (LOCAL) {
          if (mx >= best_v && rvalid[r]) {  // rare: first maximum in row-major order
            const uint32_t y = yl0 + r + 1;
            if (mx > best_v || y < best_y) { best_v = mx; best_y = y; best_x = x0 + c + 1; }
          }
        }
      }
      bottom = hup;
      if (store_d) {
        if (R == 4) *dp = (uint8_t)bits; else *reinterpret_cast<uint16_t *>(dp) = (uint16_t)bits;
      }
      if (lane == 31) rowout[s] = hup;  // bottom row of the tile, for the tile below
    }
  }
  if (lane == 0) a.corner[ti] = up_last;  // H[y0][x0+TWv] for the tile to the right
  // right edge of the tile, for the tile to the right; H[m][n] from the last tile column
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const uint32_t y = yl0 + r + 1;
    if (rvalid[r]) a.colbuf[y] = h[r];
    if (y == m && x0 + TWv == n) *a.corner_out = h[r];
  }
  if (LOCAL) {
    unsigned long long key = best_v > 0 ? wave_key(best_v, best_y, best_x) : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
      key = other > key ? other : key;
    }
    if (lane == 0 && key) atomicMax(a.best, key);
  }
}

// start cell, f, corner and status of the pair from the wavefront's reduction buffers
static __global__ void wave_finish_kernel(WaveArgs a, FillOut fo, uint32_t pair) {
  fo.corner[pair] = (double)*a.corner_out;
  if (!a.local) {
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
  }
}

}  // namespace alnr
