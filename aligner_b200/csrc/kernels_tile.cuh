// kernels_tile.cuh — intra-pair fill for ONE long pair (BASELINE config 4), second generation of
// the band pipeline of kernels_band.cuh: the same pipeline of 128-row bands (one warp per band,
// lane l owns 4 consecutive rows and runs l steps behind lane l-1), but a lane's step is a
// 4 x 4 TILE of cells instead of a 4 x 1 column:
//   * the per-step overhead (4 shuffles for the tile's top row, the query load, the D store, the
//     boundary-row hand-off) is paid once per 16 cells instead of once per 4;
//   * the 16 cells of a tile form 7 anti-diagonals, so a lane has ~2.3 independent dependency
//     chains in flight — the column kernel had one, and it was bound by single-warp latency;
//   * a tile is exactly one word of the "blocked" 2-bit D layout (4 rows x 4 columns, common.cuh):
//     one aligned 32-bit streaming store per lane per step, same traceback as before;
//   * the substitution scores of the lane's 4 rows sit in shared memory as one int4 per query
//     symbol ([symbol][lane], conflict-free 16-byte loads): 4 loads per tile instead of 16;
//   * the pipeline is 4x shorter in steps: n/4 + 32 per band instead of n + 31.
//
// Band b takes its top boundary row from band b-1 through an L2-resident row buffer WITHOUT a
// separate progress word: the buffer is preset to a sentinel that no score can take
// (0x80808080; |H| < 2^30 is guaranteed by the numeric-domain test, batch.cu::classify), lane 31 of
// the producer stores its tile's bottom row with one 16-byte st.cg, and the consumer requests the
// 32 values of its next 8 steps in the middle of the current chunk (ld.cg) and re-loads the lanes
// that still hold the sentinel.  Data and flag are the same word, so no fence and no second round
// trip.  Every band runs at the speed of the band above, so it keeps following it at the distance
// it started with: 31 steps of lane skew + one chunk + the hand-off slack (measured 47 steps,
// DESIGN.md 4.4) — the critical path of the whole fill is bands x that distance + n / 4 steps.
//
// Static penalty only, exactly like kernels_band.cuh (global mode: cell (1,1) pays del, every
// other cell ext, simple/mod.rs:72,88-92; local mode with del == ext).
#pragma once
#include <climits>

#include "kernels_band.cuh"

namespace alnr {

constexpr int kTileSentinel = (int)0x80808080u;
#ifndef ALNR_TILE_CHUNK
#define ALNR_TILE_CHUNK 8
#endif
// steps per boundary chunk: a chunk's 4 * kTileChunk boundary values are fetched by its first lanes
constexpr uint32_t kTileChunk = ALNR_TILE_CHUNK;
#ifndef ALNR_TILE_PREFETCH_AT
#define ALNR_TILE_PREFETCH_AT (ALNR_TILE_CHUNK * 5 / 8)
#endif
constexpr uint32_t kTilePrefetchAt = ALNR_TILE_PREFETCH_AT;  // step of a chunk at which the next chunk is requested

struct TileLane {
  int h[4];         // H[yl0+1+r][c0]: the column left of the lane's next tile
  int bot[4];       // bottom row of the lane's previous tile (lane+1 reads it by shuffle)
  int diag0;        // H[yl0][c0]: corner above-left of the next tile
  uint32_t qw;      // the 4 query symbols of the next tile
  const uint8_t *qp;
  uint32_t *dptr;   // D word of the next tile
  int *rowp;        // &my_row[c0] of the next tile
  int best_v;
  uint32_t best_y, best_x;
  int best_key;     // LOCAL: the band's best cell of this lane as value * 16 + (3 - r) * 4 + (3 - j) ...
  uint32_t best_ti; // ... and its tile
  int corner;
  bool has_corner;
};

// STATEFUL instantiation only (kept out of TileLane: the other instantiations keep their tuned schedule)
struct TileStateful {
  uint32_t dmw, dmw_next;     // first lane of the first band: delmap words of this and the next 32 columns
  const uint32_t *delmap;
  uint32_t *zeromap;
  bool last_band;
};

// 4 symbols starting at any byte address (two aligned words + funnel shift; the buffers carry 16
// bytes of padding, batch.cu)
__device__ __forceinline__ uint32_t tile_load_q4(const uint8_t *p) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const uint32_t *w = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
  return __funnelshift_r(__ldg(w), __ldg(w + 1), (uint32_t)(a & 3) * 8u);
}

__device__ __forceinline__ unsigned long long tile_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#ifdef ALNR_TILE_DEBUG
#define TILE_STAMP(k) do { if (lane == 0 && a.prog) reinterpret_cast<unsigned long long *>(a.prog)[(size_t)band * 8 + (k)] = tile_now(); } while (0)
#else
#define TILE_STAMP(k) do { } while (0)
#endif

__device__ __forceinline__ int tile_ld_row(const int *p) {
  int v;
  asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// One step: lane l computes tile ti = s - l (columns 4 ti + 1 .. 4 ti + 4) of its 4 rows.
// EDGE = 1 (ramp-up: the first 32 steps of a band, which every band below waits for) adds the
// ti < ntiles test and the one cell that pays `del` in global mode; EDGE = 2 (last columns and
// drain) also masks the query symbols past column n (those cells are computed and never read)
// and captures H[m][n].
template <bool LOCAL, int EDGE>
__device__ __forceinline__ void tile_step(TileLane &L, const int4 *sm_sc, const int *sm_top, uint32_t s, uint32_t lane,
                                          uint32_t yl0, uint32_t n, uint32_t m, uint32_t ntiles, int del, int ext,
                                          bool store_d, bool row_out) {
  int up[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) up[j] = __shfl_up_sync(0xffffffffu, L.bot[j], 1);
  const int4 tp = *reinterpret_cast<const int4 *>(sm_top + 4 * (s & (kTileChunk - 1)));
  if (lane == 0) { up[0] = tp.x; up[1] = tp.y; up[2] = tp.z; up[3] = tp.w; }
  const uint32_t ti = s - lane;  // wraps for s < lane
  if (EDGE && ti >= ntiles) return;
  const uint32_t c0 = ti * 4;
  const uint32_t qw = L.qw;
  L.qp += 4;
  L.qw = tile_load_q4(L.qp);
  const int corner_in = L.diag0;
  L.diag0 = up[3];
  uint32_t word = 0;
  // LOCAL: the tile's best cell as one key, value * 16 + (3 - r) * 4 + (3 - j): the first maximum in
  // row-major order (simple/mod.rs:212) is the largest key.  (Comparing every cell with the lane's
  // best and walking the 16 cells on a hit — the first form — ran the walk at nearly every step of
  // the lower-left half of a matrix of similar sequences: scores rise towards the diagonal.)
  int tkey = INT_MIN;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t sym = (qw >> (8 * j)) & 0xffu;
    if (EDGE == 2 && c0 + j >= n) sym = 0;
    const int4 s4 = sm_sc[sym * 32 + lane];
    const int sc[4] = {s4.x, s4.y, s4.z, s4.w};
    int diag = j == 0 ? corner_in : up[j - 1];
    int hup = up[j];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int pen = (EDGE && !LOCAL && r == 0 && j == 0 && yl0 == 0 && c0 == 0) ? del : ext;
      const int top = hup - pen, left = L.h[r] - pen, dg = diag + sc[r];
      const int mx = max(max(top, left), dg);
      uint32_t code = (mx == left) ? DIR_LEFT : DIR_DIAGONAL;  // enums.rs:23/:41
      code = (mx == top) ? DIR_TOP : code;                     // enums.rs:21/:39
      if (LOCAL) code = (mx == 0) ? DIR_BEGINNING : code;      // enums.rs:37
      word |= code << (8 * j + 2 * r);
      diag = L.h[r];
      L.h[r] = mx;
      hup = mx;
      if (LOCAL) {
        int k = mx * 16 + ((3 - r) * 4 + (3 - j));
        if (EDGE == 2 && c0 + j >= n) k = INT_MIN;  // columns past n are computed and never read
        tkey = max(tkey, k);
      }
    }
    L.bot[j] = hup;
    if (EDGE == 2 && c0 + j + 1 == n) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (yl0 + r + 1 == m) { L.corner = L.h[r]; L.has_corner = true; }
    }
  }
  // a later tile of the same rows only wins with a larger value or, at equal value, a smaller row
  if (LOCAL && (tkey >> 2) > (L.best_key >> 2)) { L.best_key = tkey; L.best_ti = ti; }
  if (store_d) __stcs(L.dptr, word);
  ++L.dptr;
  if (row_out) __stcg(reinterpret_cast<int4 *>(L.rowp), make_int4(L.bot[0], L.bot[1], L.bot[2], L.bot[3]));
  L.rowp += 4;
}

// tile_step for local mode with del != ext (simple/mod.rs:201-205): a cell pays del iff the cell above it
// holds 0 (it was a Beginning, and it is the cell before this one in the reference's column-major
// order); row 1 pays del where the speculated map says so (kernels_band.cuh); the last band records
// the zeros of row m.  Otherwise the code of tile_step<true, EDGE>:
// one step: lane l computes tile ti = s - l (columns 4 ti + 1 .. 4 ti + 4) of its 4 rows.
// EDGE = 1 (ramp-up: the first 32 steps of a band, which every band below waits for) adds the
// ti < ntiles test and the one cell that pays `del` in global mode; EDGE = 2 (last columns and
// drain) also masks the query symbols past column n (those cells are computed and never read)
// and captures H[m][n].
template <int EDGE>
__device__ __forceinline__ void tile_step_stateful(TileLane &L, TileStateful &Z, const int4 *sm_sc, const int *sm_top, uint32_t s, uint32_t lane,
                                          uint32_t yl0, uint32_t n, uint32_t m, uint32_t ntiles, int del, int ext,
                                          bool store_d, bool row_out) {
  int up[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) up[j] = __shfl_up_sync(0xffffffffu, L.bot[j], 1);
  const int4 tp = *reinterpret_cast<const int4 *>(sm_top + 4 * (s & (kTileChunk - 1)));
  if (lane == 0) { up[0] = tp.x; up[1] = tp.y; up[2] = tp.z; up[3] = tp.w; }
  const uint32_t ti = s - lane;  // wraps for s < lane
  if (EDGE && ti >= ntiles) return;
  const uint32_t c0 = ti * 4;
  const uint32_t qw = L.qw;
  L.qp += 4;
  L.qw = tile_load_q4(L.qp);
  const int corner_in = L.diag0;
  L.diag0 = up[3];
  uint32_t word = 0;
  int tkey = INT_MIN;  // as in tile_step
  int v[4][4];
  constexpr bool LOCAL = true;
  uint32_t dm4 = 0;  // row 1: which of the tile's 4 columns pay del
  if (yl0 == 0) {
    dm4 = Z.dmw >> (c0 & 31u);
    if ((c0 & 31u) == 28u) { Z.dmw = Z.dmw_next; Z.dmw_next = __ldg(Z.delmap + (c0 >> 5) + 2); }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t sym = (qw >> (8 * j)) & 0xffu;
    if (EDGE == 2 && c0 + j >= n) sym = 0;
    const int4 s4 = sm_sc[sym * 32 + lane];
    const int sc[4] = {s4.x, s4.y, s4.z, s4.w};
    int diag = j == 0 ? corner_in : up[j - 1];
    int hup = up[j];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      // both penalties are applied before the cell above is known to be a Beginning or not: the
      // zero test then only selects (one level less on the step's dependent chain)
      bool z = hup == 0;
      if (r == 0 && yl0 == 0) z = ((dm4 >> j) & 1u) != 0;  // H[0][x] is 0 by definition, not a Beginning
      const int top = z ? hup - del : hup - ext, left = z ? L.h[r] - del : L.h[r] - ext, dg = diag + sc[r];
      const int mx = max(max(top, left), dg);
      uint32_t code = (mx == left) ? DIR_LEFT : DIR_DIAGONAL;  // enums.rs:23/:41
      code = (mx == top) ? DIR_TOP : code;                     // enums.rs:21/:39
      if (LOCAL) code = (mx == 0) ? DIR_BEGINNING : code;      // enums.rs:37
      word |= code << (8 * j + 2 * r);
      diag = L.h[r];
      L.h[r] = mx;
      hup = mx;
      if (LOCAL) {
        int k = mx * 16 + ((3 - r) * 4 + (3 - j));
        if (EDGE == 2 && c0 + j >= n) k = INT_MIN;
        tkey = max(tkey, k);
        v[r][j] = mx;
      }
    }
    L.bot[j] = hup;
    if (EDGE == 2 && c0 + j + 1 == n) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (yl0 + r + 1 == m) { L.corner = L.h[r]; L.has_corner = true; }
    }
  }
  if ((tkey >> 2) > (L.best_key >> 2)) { L.best_key = tkey; L.best_ti = ti; }
  if (Z.last_band) {  // zeros of row m: the next pass's row 1 pays del one column later
    const uint32_t zr = m - 1u - yl0;  // the lane's row that is row m (>= 4: another lane's)
    uint32_t zm = 0;
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if ((uint32_t)r == zr && v[r][j] == 0 && c0 + j < n) zm |= 1u << j;
    if (zm) atomicOr(Z.zeromap + (c0 >> 5), zm << (c0 & 31u));
  }
  if (store_d) __stcs(L.dptr, word);
  ++L.dptr;
  if (row_out) __stcg(reinterpret_cast<int4 *>(L.rowp), make_int4(L.bot[0], L.bot[1], L.bot[2], L.bot[3]));
  L.rowp += 4;
}

// BandArgs as in kernels_band.cuh, except: rows = [nbands][npad] indexed by the 0-based column
// (npad = n rounded up to 4, preset to kTileSentinel), prog unused.
// REFILL (checkpointed traceback, wave.cu): a compile-time variant, so that the ordinary sweep keeps
// the instruction schedule it was tuned with (as run-time conditions the same lines cost it 12 %).
template <bool LOCAL, bool REFILL = false, bool STATEFUL = false>
__global__ void __launch_bounds__(32) fill_tile_kernel(BandArgs a) {
  static_assert(!STATEFUL || (LOCAL && !REFILL), "the stateful penalty only exists in local mode");
  if (STATEFUL && a.prog[0]) return;  // an earlier pass reproduced its penalty map: the results stand
  extern __shared__ int4 tsm[];  // [V][32] score quadruples, then 32 ints of boundary values
  const uint32_t lane = threadIdx.x;
  const uint32_t V = (uint32_t)a.V;
  int4 *sm_sc = tsm;
  int *sm_top = reinterpret_cast<int *>(tsm + V * 32);
  const uint32_t n = a.n, m = a.m;
  const uint32_t ntiles = (n + 3) / 4, npad = ntiles * 4;
  const int del = a.del, ext = a.ext;
  TileLane L;
  L.best_v = 0; L.best_y = 0; L.best_x = 0;
  constexpr bool refill = REFILL;  // checkpointed traceback: bands [band0, band_end), rows are final
  const uint32_t band_end = REFILL ? a.band_end : a.nbands;
  if (REFILL && a.ckpt) {  // the path crosses this group iff its current cell lies inside the group's rows
    const uint32_t cy = a.ckpt[0], cx = a.ckpt[1], done = a.ckpt[4];
    if (done || cx == 0 || cy <= a.band0 * 128u || cy > band_end * 128u) return;
  }

  for (uint32_t band = (REFILL ? a.band0 : 0u) + blockIdx.x; band < band_end; band += gridDim.x) {
    const uint32_t yl0 = band * 128u + lane * 4u;  // this lane's rows: yl0+1 .. yl0+4 (1-based)
    uint32_t trow[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const uint32_t y = yl0 + r + 1;
      const bool valid = y <= m;
      trow[r] = (valid ? (uint32_t)a.t[y - 1] : 0u) * V;
      int v = 0;
      if (!LOCAL && valid) v = (y == m) ? -(int)(m + 1) * del : -(int)y * del;  // simple/mod.rs:64-67,:70
      L.h[r] = v;
    }
    __syncwarp();
    for (uint32_t sym = 0; sym < V; ++sym) {
      int4 sc = make_int4(__ldg(a.matrix + trow[0] + sym), __ldg(a.matrix + trow[1] + sym),
                          __ldg(a.matrix + trow[2] + sym), __ldg(a.matrix + trow[3] + sym));
      if (LOCAL) {
        // rows past m (last band) must never hold the maximum: with these scores their cells stay
        // below the cell of row m above them (they reach it only through top - pen)
        constexpr int kNever = -(1 << 20);
        if (yl0 + 1 > m) sc.x = kNever;
        if (yl0 + 2 > m) sc.y = kNever;
        if (yl0 + 3 > m) sc.z = kNever;
        if (yl0 + 4 > m) sc.w = kNever;
      }
      sm_sc[sym * 32 + lane] = sc;
    }
    L.diag0 = (LOCAL || yl0 == 0) ? 0 : -(int)yl0 * del;  // H[yl0][0]
#pragma unroll
    for (int j = 0; j < 4; ++j) L.bot[j] = 0;
    L.qp = a.q;
    L.qw = tile_load_q4(L.qp);
    L.corner = 0; L.has_corner = false;
    L.best_key = 15; L.best_ti = 0;  // value 0 with the best cell code: only a positive cell wins
    const int *top_row = band ? a.rows + (size_t)(band - 1) * npad : nullptr;
    const bool store_d = yl0 < m && a.D != nullptr;  // score-only calls pass no D scratch
    const bool last_band = band + 1 == a.nbands;
    const bool row_out = lane == 31 && !last_band && !refill;
    TileStateful Z;
    if (STATEFUL) {
      Z.delmap = stateful_delmap(a); Z.zeromap = stateful_zeromap(a); Z.last_band = last_band;
      Z.dmw = Z.dmw_next = 0;
      if (yl0 == 0) { Z.dmw = __ldg(Z.delmap); Z.dmw_next = __ldg(Z.delmap + 1); }
    }
    L.dptr = a.D + (size_t)(yl0 / 4 - (REFILL ? a.strip0 : 0u)) * a.dcols;  // this lane's 4-row strip
    L.rowp = a.rows + (size_t)band * npad;
    const uint32_t steps = ntiles + 31;
    // A band cannot start before the band above has finished its first chunk, i.e. ~39 steps per
    // band after the launch.  All bands are resident from the start, so a waiting band must not
    // take issue slots from the working band on its scheduler: sleep in 1 us quanta until band
    // b - kTileWake has handed on its first tile (that happens >= kTileWake x 39 steps before this
    // band's own turn), and only then poll the own boundary row tightly.
    constexpr uint32_t kTileWake = 4;
    TILE_STAMP(0);
    if (band > kTileWake) {
      const int *far = a.rows + (size_t)(band - kTileWake) * npad;
      if (lane == 0)
        while (tile_ld_row(far) == kTileSentinel) __nanosleep(1000);
      __syncwarp();
    }
    TILE_STAMP(1);
    int pre = band ? tile_ld_row(top_row + lane) : 0;  // npad >= 4 and the buffer is padded by 32

    for (uint32_t s0 = 0; s0 < steps; s0 += kTileChunk) {
      const uint32_t col = 4 * s0 + lane;  // 0-based column of this lane's boundary value, x = col + 1
      if (4 * s0 < npad) {
        if (band) {
          // the producer's 16-byte stores land whole tiles; a lane that still sees the sentinel re-loads
          for (;;) {
            const bool need = lane < 4 * kTileChunk && col < npad && pre == kTileSentinel;
            if (!__any_sync(0xffffffffu, need)) break;
            if (need) { __nanosleep(20); pre = tile_ld_row(top_row + col); }
          }
        } else {  // H[0][x]: simple/mod.rs:59-62, :69
          const uint32_t xx = col + 1;
          pre = (LOCAL || xx > n) ? 0 : ((xx == n) ? -(int)(n + 1) * del : -(int)xx * del);
        }
        __syncwarp();
        if (s0 == 0) TILE_STAMP(2);
        if (s0 == 32) TILE_STAMP(3);
        if (s0 == 64) TILE_STAMP(4);
        if (lane < 4 * kTileChunk) sm_top[lane] = pre;
        __syncwarp();
      }
      // The next chunk's boundary values are requested in the MIDDLE of this chunk: a band that
      // follows its producer at the minimum distance finds them only if the load is issued after
      // the producer has stored them, i.e. as late as the L2 round trip allows (kTilePrefetchAt).
      const uint32_t ncol = col + 4 * kTileChunk;
      const int *pf = (band && lane < 4 * kTileChunk && ncol < npad) ? top_row + ncol : nullptr;
      pre = 0;
      // steady (EDGE = 0): every lane is inside [1, ntiles) for all steps and nobody touches column n
      const bool inner = 4 * (s0 + kTileChunk) < n;  // the chunk stays left of column n for every lane
      if (inner && s0 >= 32) {
#pragma unroll
        for (uint32_t j = 0; j < kTileChunk; ++j) {
          if (j == kTilePrefetchAt && pf) pre = tile_ld_row(pf);
          if constexpr (STATEFUL)
            tile_step_stateful<0>(L, Z, sm_sc, sm_top, s0 + j, lane, yl0, n, m, ntiles, del, ext, store_d, row_out);
          else
            tile_step<LOCAL, 0>(L, sm_sc, sm_top, s0 + j, lane, yl0, n, m, ntiles, del, ext, store_d, row_out);
        }
      } else if (inner) {
#pragma unroll
        for (uint32_t j = 0; j < kTileChunk; ++j) {
          if (j == kTilePrefetchAt && pf) pre = tile_ld_row(pf);
          if constexpr (STATEFUL)
            tile_step_stateful<1>(L, Z, sm_sc, sm_top, s0 + j, lane, yl0, n, m, ntiles, del, ext, store_d, row_out);
          else
            tile_step<LOCAL, 1>(L, sm_sc, sm_top, s0 + j, lane, yl0, n, m, ntiles, del, ext, store_d, row_out);
        }
      } else {
        const uint32_t jend = min(kTileChunk, steps - s0);
        if (pf) pre = tile_ld_row(pf);
        for (uint32_t j = 0; j < jend; ++j)
          if constexpr (STATEFUL)
            tile_step_stateful<2>(L, Z, sm_sc, sm_top, s0 + j, lane, yl0, n, m, ntiles, del, ext, store_d, row_out);
          else
            tile_step<LOCAL, 2>(L, sm_sc, sm_top, s0 + j, lane, yl0, n, m, ntiles, del, ext, store_d, row_out);
      }
    }
    TILE_STAMP(5);
    if (LOCAL && (L.best_key >> 4) > 0) {  // the band's best cell of this lane into the running arg-max
      const int v = L.best_key >> 4;
      const uint32_t y = yl0 + (3u - (((uint32_t)L.best_key >> 2) & 3u)) + 1u;
      const uint32_t x = 4u * L.best_ti + (3u - ((uint32_t)L.best_key & 3u)) + 1u;
      if (v > L.best_v || (v == L.best_v && y < L.best_y)) { L.best_v = v; L.best_y = y; L.best_x = x; }
    }
    if (last_band && L.has_corner && !refill) *a.corner_out = L.corner;
    __syncwarp();
  }
  if (LOCAL && !refill) {
    unsigned long long key = L.best_v > 0 ? band_key(L.best_v, L.best_y, L.best_x) : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
      key = other > key ? other : key;
    }
    if (lane == 0 && key) atomicMax(a.best, key);
  }
}

}  // namespace alnr
