// kernels_band.cuh — intra-pair fill for ONE long pair (BASELINE config 4: ~10^5 x 10^5 global
// alignment, 10^10 cells) as a single persistent launch: a pipeline of row bands.
//
// The matrix is cut into bands of 32 x R target rows (R = 4 rows per lane; R = 2 is instantiated
// and measured slower).  One WARP owns a band and sweeps it over all columns: lane l keeps its R
// consecutive rows in registers and runs l columns behind lane l-1 (the bottom cell of lane l-1
// arrives by one __shfl_up per column).  Band b takes its top boundary
// row from band b-1 through an L2-resident row buffer: the producer publishes "columns done"
// every 32 columns (threadfence + volatile store), the consumer polls it before loading the next
// 32 boundary values.  All bands are resident at once (one 32-thread block per band, spread over
// the 148 SMs x 4 schedulers; more bands than resident warps are taken in increasing order by
// the same warps), so the dependency chain always points at a running warp — no per-diagonal
// launches, no inter-kernel tails.  The anti-diagonal wavefront of the classic formulation is
// what this pipeline looks like in flight.
//
// Directions go to the "blocked" 2-bit D layout (common.cuh): a lane packs R rows x 16/R columns
// into one aligned 32-bit streaming store (0.25 B/cell; the only HBM stream of the kernel).
//
// Measured (B200, scripts/band_probe.py): one warp alone needs ~280 cycles per step (~70 dependent
// instructions at ~4 cycles each), i.e. the kernel is bound by single-warp instruction latency and
// by the pipeline fill (bands x 64 steps), not by issue slots or HBM.
//
// Valid for a STATIC penalty only: global mode (cell (1,1) pays del, every other cell ext —
// get_direction never returns Beginning, simple/mod.rs:88-92) and local mode with del == ext.
// Local mode with del != ext carries the penalty from cell (m, x-1) to (1, x)
// (simple/mod.rs:201-205), which serialises columns: such pairs use the warp kernel.
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

__device__ __forceinline__ unsigned long long band_key(int v, uint32_t y, uint32_t x) {
  // larger value first, then smaller y, then smaller x (row-major first maximum, simple/mod.rs:212)
  return ((unsigned long long)(uint32_t)v << 40) | ((unsigned long long)(0xFFFFFu - y) << 20) |
         (unsigned long long)(0xFFFFFu - x);
}

static __global__ void band_init_kernel(BandArgs a) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < a.nbands) a.prog[i] = 0;
  if (i == 0) { *a.best = 0ull; *a.corner_out = 0; }
}

// Per-lane state of one band sweep (registers).  Every address is a running pointer so that a
// step is pure register arithmetic plus its loads and stores.
template <int R>
struct BandLane {
  uint32_t soff[R];     // shared-memory byte offset of the matrix row of each of the 4 target symbols
  int h[R];             // H[y][x-1] of the lane's rows
  int diag0;            // H[yl0][x-1]
  int bottom;           // H[yl0+4][x] of the previous step (handed to lane+1)
  int rowchunk;         // lane j: top boundary value of column chunk_base + j + 1
  uint32_t qn;          // query symbol of the lane's next column
  uint32_t dacc;        // 4 columns x 4 rows of directions being packed
  const uint8_t *qp;    // &q[x] of the lane's next column + 1
  uint32_t *dptr;       // D word being packed
  int *rowp;            // &my_row[x] of the lane's next column
  int best_v;
  uint32_t best_y, best_x;
};

// One step of the skewed sweep: lane l computes column c = s - l of its 4 rows.
// EDGE = true adds what only the ramp-up / drain steps need: the c < n test and the one cell that
// pays `del` in global mode, (1, 1) (simple/mod.rs:72,88-92).
template <int R, bool LOCAL, bool EDGE>
__device__ __forceinline__ void band_step(BandLane<R> &L, const char *smem, uint32_t s, uint32_t lane, uint32_t yl0,
                                          uint32_t n, uint32_t m, int del, int ext, bool store_d, bool row_out) {
  int up = __shfl_up_sync(0xffffffffu, L.bottom, 1);
  const int top0 = __shfl_sync(0xffffffffu, L.rowchunk, s & 31);
  if (lane == 0) up = top0;
  const uint32_t c = s - lane;  // 0-based column; wraps for s < lane
  if (EDGE && c >= n) return;
  const uint32_t qs4 = L.qn * 4;
  L.qn = *L.qp++;  // one past the end at x == n: the buffer is padded, the value is never used
  int diag = L.diag0;
  L.diag0 = up;
  int hup = up;
  uint32_t bits = 0;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int pen = (EDGE && !LOCAL && r == 0 && yl0 == 0 && c == 0) ? del : ext;
    const int sc = *reinterpret_cast<const int *>(smem + L.soff[r] + qs4);
    const int top = hup - pen, left = L.h[r] - pen, dg = diag + sc;
    const int mx = max(max(top, left), dg);
    uint32_t code = (mx == left) ? DIR_LEFT : DIR_DIAGONAL;  // enums.rs:23/:41
    code = (mx == top) ? DIR_TOP : code;                     // enums.rs:21/:39
    if (LOCAL) code = (mx == 0) ? DIR_BEGINNING : code;      // enums.rs:37
    bits |= code << (2 * r);
    diag = L.h[r];
    L.h[r] = mx;
    hup = mx;
    if (LOCAL) {
      if (mx >= L.best_v) {  // rare: first maximum in row-major order (simple/mod.rs:212)
        const uint32_t y = yl0 + r + 1;
        if (y <= m && (mx > L.best_v || y < L.best_y)) { L.best_v = mx; L.best_y = y; L.best_x = c + 1; }
      }
    }
  }
  L.bottom = hup;
  constexpr uint32_t C = 16 / R;  // columns per D word
  L.dacc |= bits << (2 * R * (c & (C - 1)));
  if ((c & (C - 1)) == C - 1 || (EDGE && c + 1 == n)) {
    if (store_d) __stcs(L.dptr, L.dacc);
    ++L.dptr;
    L.dacc = 0;
  }
  if (row_out) __stcg(L.rowp, hup);
  ++L.rowp;
}

template <int R, bool LOCAL>
__global__ void __launch_bounds__(32) fill_band_kernel(BandArgs a) {
  constexpr uint32_t kBandRows = 32 * R;
  extern __shared__ int bsm[];  // V*V
  const uint32_t lane = threadIdx.x;
  for (int i = lane; i < a.V * a.V; i += 32) bsm[i] = a.matrix[i];
  __syncwarp();
  const uint32_t n = a.n, m = a.m;
  const int del = a.del, ext = a.ext;
  const char *smem = reinterpret_cast<const char *>(bsm);
  BandLane<R> L;
  L.best_v = 0; L.best_y = 0; L.best_x = 0;

  for (uint32_t band = blockIdx.x; band < a.nbands; band += gridDim.x) {
    const uint32_t yl0 = band * kBandRows + lane * R;  // this lane's rows: yl0+1 .. yl0+R (1-based)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t y = yl0 + r + 1;
      const bool valid = y <= m;
      L.soff[r] = (valid ? (uint32_t)a.t[y - 1] : 0u) * (uint32_t)a.V * 4u;
      int v = 0;
      if (!LOCAL && valid) v = (y == m) ? -(int)(m + 1) * del : -(int)y * del;  // simple/mod.rs:64-67,:70
      L.h[r] = v;
    }
    L.diag0 = (LOCAL || yl0 == 0) ? 0 : -(int)yl0 * del;  // H[yl0][0]
    L.bottom = 0; L.rowchunk = 0; L.dacc = 0;
    L.qn = a.q[0];
    L.qp = a.q + 1;
    const int *top_row = band ? a.rows + (size_t)(band - 1) * (n + 1) : nullptr;
    int *my_row = a.rows + (size_t)band * (n + 1);
    volatile uint32_t *prog_prev = band ? a.prog + (band - 1) : nullptr;
    volatile uint32_t *prog_mine = a.prog + band;
    const bool store_d = yl0 < m && a.D != nullptr;  // score-only calls pass no D scratch
    const bool last_band = band + 1 == a.nbands;
    const bool row_out = lane == 31 && !last_band;
    L.dptr = a.D + (size_t)(yl0 / R) * a.dcols;  // this lane's R-row strip
    L.rowp = my_row + 1;
    const uint32_t steps = n + 31;

    // chunks of 32 steps: the boundary values of the chunk are fetched at its start, lane 31
    // finishes column s0 (a multiple of 32) at step s0 + 30 and publishes it there
    for (uint32_t s0 = 0; s0 < steps; s0 += 32) {
      if (s0 < n) {
        const uint32_t xx = s0 + lane + 1;
        if (band) {
          // wait until the band above has published the next 32 values of its bottom row
          const uint32_t need = min(n, s0 + 32);
          if (lane == 0) { while (*prog_prev < need) __nanosleep(32); __threadfence(); }
          __syncwarp();
          L.rowchunk = xx <= n ? __ldcg(top_row + xx) : 0;
        } else {  // H[0][x]: simple/mod.rs:59-62, :69
          L.rowchunk = (LOCAL || xx > n) ? 0 : ((xx == n) ? -(int)(n + 1) * del : -(int)xx * del);
        }
      }
      const bool steady = s0 >= 32 && s0 + 32 <= n;  // every lane is inside [0, n) for all 32 steps
      if (steady) {
#pragma unroll 4
        for (uint32_t j = 0; j < 32; ++j)
          band_step<R, LOCAL, false>(L, smem, s0 + j, lane, yl0, n, m, del, ext, store_d, row_out);
      } else {
        const uint32_t jend = min(32u, steps - s0);
        for (uint32_t j = 0; j < jend; ++j)
          band_step<R, LOCAL, true>(L, smem, s0 + j, lane, yl0, n, m, del, ext, store_d, row_out);
      }
      // lane 31 has now finished every column up to min(n, s0 + 1): publish that for the band below
      // (its chunk s0' needs s0' + 32 columns, so it runs two chunks behind)
      if (row_out && s0 >= 32) { __threadfence(); *prog_mine = min(n, s0 + 1); }
    }
    if (last_band) {
#pragma unroll
      for (int r = 0; r < R; ++r) if (yl0 + r + 1 == m) *a.corner_out = L.h[r];
    }
    __syncwarp();
  }
  if (LOCAL) {
    unsigned long long key = L.best_v > 0 ? band_key(L.best_v, L.best_y, L.best_x) : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
      key = other > key ? other : key;
    }
    if (lane == 0 && key) atomicMax(a.best, key);
  }
}

// start cell, f, corner and status of the pair from the reduction buffers
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
  }
}

}  // namespace alnr
