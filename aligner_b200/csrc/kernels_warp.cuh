// kernels_warp.cuh — intra-pair fill with ONE WARP PER PAIR (or per (pair, matrix) item).
//
// Serves everything between the short-read strip kernel and the 10^5-scale wavefront kernel:
// single-pair calls (BASELINE configs 1, 2 and the config-4 file), protein / ragged batches in
// exact INT32, and the FP64 heuristic matrix search (config 5), where every item has its own
// non-integral matrix (the warp stages it in shared memory; one thread per item would have to
// look 4.6 KB tables up in L2).
//
// Lane l owns R consecutive target rows in registers and works l columns behind lane l-1, so the
// bottom cell of lane l-1 arrives by one __shfl_up per column (skewed systolic sweep, 31 bubble
// steps per band).  Targets longer than 32*R rows are cut into bands; a band's bottom row
// passes to the next through an L2-resident row buffer.  The words of 2-bit directions a warp
// emits in one step are consecutive in the "skewed" D layout (common.cuh): 128 bytes per store.
//
// The recurrence is the reference's, literally (simple/mod.rs:74-97 global, :187-210 local;
// enums.rs:18-46): top/left pay the carried `penalty`, ties resolve Top > Left > Diagonal with
// |max - cand| < EPSILON, Beginning iff max == 0, penalty <- ext unless the cell just computed
// is Beginning.  Inside a column the penalty follows the cell above, across lanes it follows the
// value handed up by the shuffle (Beginning <=> value == 0 for interior cells).  The only
// cross-column dependency — penalty(1, x) = del iff H[m][x-1] == 0 (local mode, del != ext) — is
// speculated exactly like the strip kernel: assume a bit vector z, observe the true bits on
// row m, repeat the pass with the observed bits until they agree (each pass extends the verified
// prefix); pairs that do not settle within kWarpPasses go to the exact column-major kernel.
#pragma once
#include "common.cuh"
#include "kernels_exact.cuh"
#include "warp_args.h"

namespace alnr {

__device__ __forceinline__ int shfl_up_t(int v) { return __shfl_up_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ double shfl_up_t(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }

// grid: ceil(count / kWarpsPerBlock); smem: V*V*sizeof(T) * (item_matrices ? kWarpsPerBlock : 1)
// max of three for the warp kernel.  FP64: the host (shared matrix) or the staging loop below
// (per-item matrices) guarantees finite inputs far from overflow, so no NaN can arise and
// f64::max (enums.rs:19) reduces to compare-select — 1 DSETP + 2 selects instead of fmax's
// DSETP + selects + NaN fix-up.  (Only the sign of a zero result can differ; `== 0.0` and the
// EPSILON tests do not see it.)
__device__ __forceinline__ int wmax3(int a, int b, int c) { return max(max(a, b), c); }
// score lookups by 32-bit shared address: row base per DP row, one add per cell
__device__ __forceinline__ int lds_score(uint32_t addr, int) {
  int v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ double lds_score(uint32_t addr, double) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ double wmax3(double a, double b, double c) {
  const double g = a > b ? a : b;
  return g > c ? g : c;
}
constexpr double kWarpFiniteBound = 1e290;

// |max - cand| < f64::EPSILON (enums.rs:21,23,39,41) when max >= cand is known — here max was picked
// from the candidates by compare-select over finite values, so max - cand >= +0.0 exactly and the
// fabs is dropped.  (Measured alternatives, profiles/r02_fill_warp_f64_ncu.txt: the alu pipe, with
// the selects, is this kernel's busiest at 68 %, the FP64 pipe has room — so the test stays a DSETP;
// as an integer compare of the difference's high word against 0x3CB00000 it cost two alu
// instructions per cell and gained nothing.  Predicated IMAD moves instead of FSEL for the max are
// if-converted back into selects by ptxas.)  The exact kernel keeps the literal fabs / fmax form.
__device__ __forceinline__ bool same_ge(int mx, int c) { return mx == c; }
__device__ __forceinline__ bool same_ge(double mx, double c) { return __dsub_rn(mx, c) < DBL_EPSILON; }
// "this cell may reach the running maximum": a conservative test that only triggers the ordered
// rescan (which compares exactly).  FP64: the running maximum is never negative, so mx >= best
// implies hi32(mx) >= hi32(best) as signed integers (a negative mx has a negative high word) — one
// ISETP per cell.  Written as `mx >= best`, nvcc turns the OR over a lane's rows into an fmax chain
// (DSETP.MAX + 2 selects + NaN fix-up + moves, ~8 instructions per cell).
__device__ __forceinline__ bool may_hit(int mx, int best) { return mx >= best; }
__device__ __forceinline__ bool may_hit(double mx, double best) {
  return __double2hiint(mx) >= __double2hiint(best);
}

// MODE: 0 global; 1 local with del == ext (static penalty: saves two selects per cell); 2 local, any
// del / ext (a.stateful says whether the speculation is needed).
// History: an intermediate round-1 build (1db19ef) saw wrong scores from the INT32 instantiation of
// MODE 1 and routed INT32 around it.  The two places where MODE 1 and MODE 2 differed in generated
// code then — the inline per-cell arg-max update and the rows below row m — were rewritten soon
// after (a5e2e93: deferred rescan, idle lanes), and since then the failure does not reproduce: every
// R = 2..16, INT32 and FP64, and the whole parity suite are bit-exact in BOTH modes on the B200
// (compute-sanitizer is closed on this pool).  MODE 1 serves INT32 again;
// ALNR_WARP_INT_MODE2=1 restores the old routing.
template <typename T, int R, int MODE, bool MATS>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) fill_warp_kernel(WarpArgs<T> a) {
  constexpr bool LOCAL = MODE != 0;
  extern __shared__ unsigned char smem_raw[];
  T *smem = reinterpret_cast<T *>(smem_raw);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // PWM view (common.cuh): the shared matrix is the 4 x L profile, a row of it per target symbol
  const bool pwm = a.cv.pwm_len != 0;
  const int V = pwm ? (int)a.cv.pwm_len : a.V, VV = pwm ? 4 * V : V * V;
  uint32_t slot = blockIdx.x * kWarpsPerBlock + warp;
  bool have = slot < a.cv.count;
  if (a.list) {  // second launch of the incremental speculation: the slots the verify kernel left
    have = slot < *a.list_count;
    if (have) slot = a.list[slot];
  }
  uint32_t pair = 0;
  if (have) {
    pair = a.cv.order ? a.cv.order[a.cv.first + slot] : a.cv.first + slot;
    // rejected by validation; 11 may be left over from the first launch
    if (a.fo.status[pair] != 0 && !(a.list && a.fo.status[pair] == 11)) have = false;
  }
  const T *smat;
  if (a.item_matrices) {
    T *mine = smem + (size_t)warp * VV;
    bool bad = false;
    if (have) for (int i = lane; i < VV; i += 32) {
      const T v = a.item_matrices[(size_t)pair * VV + i];
      mine[i] = v;
      if (sizeof(T) == 8) bad |= !(fabs((double)v) < kWarpFiniteBound);
    }
    __syncwarp();
    smat = mine;
    if (sizeof(T) == 8 && __any_sync(0xffffffffu, bad)) {
      // non-finite matrix entry: the exact kernel keeps the reference's NaN-ignoring max
      if (have && lane == 0) a.fail_list[atomicAdd(a.fail_count, 1u)] = slot;
      have = false;
    }
  } else {
    for (int i = threadIdx.x; i < VV; i += blockDim.x) smem[i] = a.matrix[i];
    __syncthreads();
    smat = smem;
  }
  if (!have) return;
  const uint32_t smat_addr = (uint32_t)__cvta_generic_to_shared(smat);

  const uint32_t sp = a.cv.seq(pair);
  const uint8_t *q, *t;
  uint32_t n, m;
  a.cv.pair_dims(sp, q, t, n, m);
  constexpr uint32_t BH = 32 * R;  // rows per band
  const uint32_t nbands = (m + BH - 1) / BH;
  const uint32_t m_lane = ((m - 1) % BH) / R, m_r = (m - 1) % R;  // where row m lives (last band)
  const T del = a.del, ext = a.ext;
  const T ndel0 = nsub(T(0), del);  // 0 - del: "top" below a Beginning cell
  const bool stateful = MODE == 2 && a.stateful;  // MODE 2 also runs static-penalty problems
  T *rowbuf = a.rowbuf + (size_t)slot * a.rowbuf_stride;
  uint32_t *Dslot = a.D + (uint64_t)slot * a.dl.nstrips * (a.dl.ncols + 31);
  const uint64_t band_words = (uint64_t)(a.dl.ncols + 31) * 32;
  const size_t W = (size_t)n + 1;
  uint32_t *zb0 = stateful ? a.zbuf + (size_t)slot * 2 * a.nzw : nullptr;
  // first launch: assume all-zero bits, observe into zb0; second launch: start from those
  const uint32_t *zin = (stateful && a.list) ? zb0 : nullptr;
  uint32_t *zout = (stateful && a.list) ? zb0 + a.nzw : zb0;
  // one pass only, mispredicted slots go to the verify kernel (needs lane 0's bottom row: rowbuf is
  // free when there is a single band)
  const bool defer = stateful && !MATS && a.vlist && !a.list && nbands == 1 && m > (uint32_t)R;
  if (a.list && lane == 0) a.fo.status[pair] = 0;

  // row 0 and column 0 of the optional full matrices (simple/mod.rs:59-70; zeros/Beginning local)
  if (MATS) {
    for (uint32_t x = lane; x <= n; x += 32) {
      T v = T(0);
      if (!LOCAL && x > 0) v = (x == n) ? -(T(n) + T(1)) * del : -T(x) * del;
      a.H_out[x] = (double)v;
      a.D_out[x] = LOCAL ? DIR_BEGINNING : (x == 0 ? DIR_BEGINNING : DIR_LEFT);
    }
    for (uint32_t y = 1 + lane; y <= m; y += 32) {
      T v = T(0);
      if (!LOCAL) v = (y == m) ? -(T(m) + T(1)) * del : -T(y) * del;
      a.H_out[y * W] = (double)v;
      a.D_out[y * W] = LOCAL ? DIR_BEGINNING : DIR_TOP;
    }
  }

  if (m == 0) {  // only PWM windows get here (pwm/mod.rs:76-81: arg-max (0,0), empty alignment)
    if (lane == 0) {
      a.fo.f[pair] = 0.0; a.fo.corner[pair] = 0.0;
      a.fo.start[2 * (size_t)pair] = 0; a.fo.start[2 * (size_t)pair + 1] = 0;
    }
    return;
  }
  T best = T(0), corner = T(0);
  uint32_t by = 0, bx = 0;
  bool settled = true;
  for (int pass = 0;; ++pass) {
    best = T(0); by = 0; bx = 0;
    bool mism = false;
    for (uint32_t band = 0; band < nbands; ++band) {
      const uint32_t yb = band * BH + lane * R;  // this lane's rows: yb+1 .. yb+R (1-based)
      const bool last_band = band + 1 == nbands;
      // The lane's R scores ping-pong between two register arrays (two steps per loop trip): the old
      // H[y][x-1] — still needed as the diagonal of row y+1 — and the new H[y][x] never compete for a
      // register, so the step loop carries no register moves (they were ~5 of 30 instructions per
      // FP64 cell).  Both arrays start as column 0: a lane idles until step `lane`.
      uint32_t toff[R];
      T hA[R], hB[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const uint32_t y = yb + r + 1;
        const bool valid = y <= m;
        toff[r] = smat_addr + (valid ? (uint32_t)t[y - 1] * V : 0u) * (uint32_t)sizeof(T);  // row of the matrix
        T v = T(0);
        if (!LOCAL && valid) v = (y == m) ? -(T(m) + T(1)) * del : -T(y) * del;  // :64-67, :70
        hA[r] = v;
        hB[r] = v;
      }
      T dprev = (LOCAL || yb == 0) ? T(0) : -T(yb) * del;  // H[yb][0]
      T bottom = T(0);
      uint32_t zw = 0, zacc = 0;
      uint32_t qn = pwm ? 0u : q[0];
      uint32_t *dp = Dslot + band * band_words + lane;
      const bool store_d = yb < m && a.D != nullptr;  // score-only calls pass no D scratch
      const bool obs_lane = last_band && lane == m_lane;
      const bool top_lane = band == 0 && lane == 0;
      const uint32_t steps = n + 31;
      auto step = [&](const T (&hin)[R], T (&h)[R], const uint32_t s) {
        T up = shfl_up_t(bottom);
        const uint32_t c = s - lane;  // 0-based column; wraps for s < lane
        if (c < n && yb < m) {  // lanes whose rows all lie below row m idle (last band)
          const uint32_t x = c + 1;
          if (lane == 0) {
            if (band == 0) { up = T(0); if (!LOCAL) up = (x == n) ? -(T(n) + T(1)) * del : -T(x) * del; }  // :59-62, :69
            else up = rowbuf[x];
          }
          const uint32_t qoff = qn * (uint32_t)sizeof(T);  // PWM view: the column index itself (pwm/mod.rs:62)
          if (x < n) qn = pwm ? x : q[x];
          T diag = dprev;
          dprev = up;
          T hup = up;
          // penalty of the lane's first row (simple/mod.rs:72,88-92 / :185,201-205)
          T pen = ext;
          if (!LOCAL) {
            if (top_lane && c == 0) pen = del;           // the very first cell pays del
          } else if (top_lane) {
            bool z = c == 0;                              // initial penalty = del
            if (stateful && c > 0) {
              if (((c - 1) & 31) == 0) zw = zin ? zin[(c - 1) >> 5] : 0u;
              z = (zw >> ((c - 1) & 31)) & 1u;            // assumed: H[m][x-1] == 0
            }
            pen = z ? del : ext;
          } else {
            pen = Num<T>::is_zero(up) ? del : ext;        // cell above is Beginning iff it is 0
          }
          uint32_t bits = 0;
          bool hit = false;  // some cell of this step reaches the running maximum (rare)
          bool zp = false;   // MODE 2: the cell above is Beginning
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const T sc = lds_score(toff[r] + qoff, T(0));
            T top, lft;
            if (MODE == 2 && r > 0) {
              // penalty = del iff the cell above is Beginning, i.e. hup == 0: both candidates are
              // formed off the dependent chain, the zero test of the cell above only selects
              const T le = nsub(hin[r], ext), ld = nsub(hin[r], del);
              top = zp ? ndel0 : nsub(hup, ext);
              lft = zp ? ld : le;
            } else {
              top = nsub(hup, pen); lft = nsub(hin[r], pen);
            }
            const T dg = nadd(diag, sc);
            const T mx = wmax3(top, lft, dg);
            uint32_t d = same_ge(mx, lft) ? DIR_LEFT : DIR_DIAGONAL;  // enums.rs:23/:41
            d = same_ge(mx, top) ? DIR_TOP : d;                       // enums.rs:21/:39
            if (LOCAL) {
              const bool zero = Num<T>::is_zero(mx);                       // enums.rs:37
              d = zero ? DIR_BEGINNING : d;
              if (MODE == 2) zp = zero;
            }
            if (MODE != 2) pen = ext;
            bits |= d << (2 * r);
            diag = hin[r];
            h[r] = mx;
            hup = mx;
            if (LOCAL) hit |= may_hit(mx, best);
            if (MATS) {
              const uint32_t y = yb + r + 1;
              if (y <= m) { a.H_out[y * W + x] = (double)mx; a.D_out[y * W + x] = (uint8_t)d; }
            }
          }
          if (LOCAL && hit) {
            // first maximum in row-major order (simple/mod.rs:212): the rows of this step in
            // ascending order, exactly the updates a per-cell test would make
#pragma unroll
            for (int r = 0; r < R; ++r) {
              const uint32_t y = yb + r + 1;
              const T v = h[r];
              if (y <= m && (v > best || (v == best && y < by))) { best = v; by = y; bx = x; }
            }
          }
          bottom = hup;
          if (store_d) *dp = bits;
          if (lane == 31 && !last_band) rowbuf[x] = hup;
          if (defer && lane == 0) rowbuf[x] = hup;  // row R, kept for warp_verify_kernel
          if (obs_lane && (stateful || x == n)) {
            T hm = h[0];
#pragma unroll
            for (int r = 1; r < R; ++r) if ((uint32_t)r == m_r) hm = h[r];
            if (x == n) corner = hm;
            if (stateful) {
              // observe z[x-1] = (H[m][x] == 0): it sets the penalty of cell (1, x+1)
              zacc |= (Num<T>::is_zero(hm) ? 1u : 0u) << (c & 31);
              if ((c & 31) == 31 || x == n) {
                const uint32_t w = c >> 5;
                const uint32_t assumed = zin ? zin[w] : 0u;
                // bits at column index >= n-1 never feed a penalty (there is no column n+1)
                const uint32_t nb = (n - 1) - (w << 5);
                const uint32_t vmask = (n - 1 <= (w << 5)) ? 0u : (nb >= 32 ? 0xffffffffu : ((1u << nb) - 1u));
                if ((zacc ^ assumed) & vmask) mism = true;
                zout[w] = zacc;
                zacc = 0;
              }
            }
          }
        }
        dp += 32;
      };
      uint32_t s = 0;
#pragma unroll 1
      for (; s + 1 < steps; s += 2) {
        step(hA, hB, s);
        step(hB, hA, s + 1);
      }
      if (s < steps) step(hA, hB, s);
      __syncwarp();  // rowbuf / z words written by one lane are read by another
    }
    if (!stateful) break;
    if (!__any_sync(0xffffffffu, mism)) break;
    if (defer) {  // provisional results are written below; the verify kernel takes it from there
      if (lane == 0) a.vlist[atomicAdd(a.vcount, 1u)] = slot;
      break;
    }
    if (pass + 1 == kWarpPasses) { settled = false; break; }
    zin = zout;
    zout = (zout == zb0) ? zb0 + a.nzw : zb0;
  }
  if (!settled) {
    if (lane == 0) a.fail_list[atomicAdd(a.fail_count, 1u)] = slot;
    return;
  }

  // ---- per-pair results -------------------------------------------------------------------
  corner = __shfl_sync(0xffffffffu, corner, m_lane);
  if (LOCAL) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const T ov = __shfl_xor_sync(0xffffffffu, best, o);
      const uint32_t oy = __shfl_xor_sync(0xffffffffu, by, o), ox = __shfl_xor_sync(0xffffffffu, bx, o);
      // larger value, then smaller y, then smaller x; (0, 0) = "no positive cell" loses to any cell
      const bool take = ov > best || (ov == best && oy != 0 && (by == 0 || oy < by || (oy == by && ox < bx)));
      if (take) { best = ov; by = oy; bx = ox; }
    }
  }
  if (lane == 0) {
    a.fo.corner[pair] = (double)corner;
    if (LOCAL) {
      a.fo.f[pair] = (double)best;  // simple/mod.rs:247
      a.fo.start[2 * (size_t)pair] = by;
      a.fo.start[2 * (size_t)pair + 1] = bx;
      // ALNR_NO_POSITIVE_CELL (simple/mod.rs:214); the PWM aligner returns an empty alignment
      if ((by == 0 || bx == 0) && !pwm) a.fo.status[pair] = 11;
    } else {
      a.fo.f[pair] = 0.0;  // simple/mod.rs:139
      a.fo.start[2 * (size_t)pair] = m;
      a.fo.start[2 * (size_t)pair + 1] = n;
    }
  }
}

// Incremental speculation, step 2 (see WarpArgs::vlist).  One THREAD per mispredicted slot recomputes
// the rows lane 0 held (1..R) over all columns with the observed bits, in the reference's own order
// (x outer, y inner; only row 1 takes its penalty from the bits).  A wrong penalty disturbs the
// matrix only a few rows deep, so row R almost always comes out as the first launch kept it in
// rowbuf: then every row below, row m and the observed bits stand, the direction words of lane 0
// are rewritten and the arg-max is re-derived from the best P of rows 1..R and the previous arg-max
// G (G below row R is still the first maximum of the unchanged rows: P wins iff P.v >= G.v; G
// inside rows 1..R stands only if P.v >= G.v, else undecided).  Changed or undecided slots go on
// rlist for the second launch of the warp kernel.
template <typename T>
__global__ void __launch_bounds__(128) warp_verify_kernel(WarpArgs<T> a) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= *a.vcount) return;
  const uint32_t slot = a.vlist[i];
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first + slot] : a.cv.first + slot;
  const bool pwm = a.cv.pwm_len != 0;
  const int stride = pwm ? (int)a.cv.pwm_len : a.V, elems = pwm ? 4 * stride : a.V * a.V;
  const T *mat = a.item_matrices ? a.item_matrices + (size_t)pair * elems : a.matrix;
  const uint32_t sp = a.cv.seq(pair);
  const uint8_t *q, *t;
  uint32_t n, m;
  a.cv.pair_dims(sp, q, t, n, m);
  const uint32_t R = a.dl.rows_per_word;  // m > R and a single band: checked by the first launch
  const T del = a.del, ext = a.ext;
  const uint32_t *zobs = a.zbuf + (size_t)slot * 2 * a.nzw;
  const T *kept = a.rowbuf + (size_t)slot * a.rowbuf_stride;
  uint32_t *Dslot = a.D ? a.D + (uint64_t)slot * a.dl.nstrips * (a.dl.ncols + 31) : nullptr;
  constexpr int kMaxR = 16;
  T h[kMaxR];
  uint32_t toff[kMaxR];
#pragma unroll
  for (int r = 0; r < kMaxR; ++r) {
    h[r] = T(0);
    toff[r] = (uint32_t)r < R ? (uint32_t)t[r] * stride : 0u;
  }
  T best = T(0);
  uint32_t by = 0, bx = 0, zw = 0;
  bool changed = false;
  for (uint32_t x = 1; x <= n; ++x) {
    const uint32_t qs = pwm ? x - 1 : q[x - 1];
    bool z = x == 1;  // initial penalty = del (simple/mod.rs:185)
    if (x > 1) {
      if (((x - 2) & 31) == 0) zw = zobs[(x - 2) >> 5];
      z = (zw >> ((x - 2) & 31)) & 1u;  // H[m][x-1] == 0
    }
    T pen = z ? del : ext;
    T diag = T(0), hup = T(0);
    uint32_t bits = 0;
#pragma unroll
    for (int r = 0; r < kMaxR; ++r) {
      if ((uint32_t)r < R) {
        const T sc = mat[toff[r] + qs];
        const T top = nsub(hup, pen), lft = nsub(h[r], pen), dg = nadd(diag, sc);
        const T mx = wmax3(top, lft, dg);
        uint32_t d = Num<T>::same(mx, lft) ? DIR_LEFT : DIR_DIAGONAL;
        d = Num<T>::same(mx, top) ? DIR_TOP : d;
        const bool zero = Num<T>::is_zero(mx);
        d = zero ? DIR_BEGINNING : d;
        pen = zero ? del : ext;
        bits |= d << (2 * r);
        diag = h[r];
        h[r] = mx;
        hup = mx;
        const uint32_t y = r + 1;
        if (mx > best || (mx == best && y < by)) { best = mx; by = y; bx = x; }
      }
    }
    changed |= !(hup == kept[x]);
    if (Dslot) Dslot[(uint64_t)(x - 1) * 32] = bits;  // band 0, lane 0 of the skewed layout
  }
  bool decided = !changed;
  if (decided) {
    const T gv = (T)a.fo.f[pair];
    const uint32_t gy = a.fo.start[2 * (size_t)pair];
    const bool p_ok = by != 0 && bx != 0;
    if (gy == 0 || (p_ok && best >= gv)) {
      a.fo.f[pair] = (double)best;
      a.fo.start[2 * (size_t)pair] = by;
      a.fo.start[2 * (size_t)pair + 1] = bx;
      a.fo.status[pair] = (p_ok || pwm) ? 0 : 11;  // ALNR_NO_POSITIVE_CELL (simple/mod.rs:214)
    } else if (gy <= R) {
      decided = false;
    }
  }
  if (!decided) a.rlist[atomicAdd(a.rcount, 1u)] = slot;
}

}  // namespace alnr
