// kernels_diff.cuh — intra-pair fill for ONE long pair in GLOBAL mode on the DIFFERENCE recurrence
// (third generation of the long-pair path, BASELINE config 4; simple/mod.rs:42-144).
//
// The tile pipeline (kernels_tile.cuh) carries 32-bit scores and spends ~10.6 instructions per cell
// on one dependent chain per lane.  Scores of a long pair do not fit 16 bits, but their DIFFERENCES
// do.  With a = H[y-1][x-1], dh_up = H[y-1][x] - a, dv_left = H[y][x-1] - a the reference's cell
//   H[y][x] = max(H[y-1][x] - p, H[y][x-1] - p, a + s)                      simple/mod.rs:82-86
// becomes            z = H[y][x] - a = max(dh_up - p, dv_left - p, s)
//                    dv[y][x] = H[y][x] - H[y-1][x] = z - dh_up,   dh[y][x] = z - dv_left
// — exact integer arithmetic, so the directions (and with them every output) are bit-identical.
// Biased by B = 2 del (every boundary difference is >= -2 del: the -(n+1) del / -(m+1) del corner
// quirks of simple/mod.rs:69-70 included) all quantities are non-negative 16-bit values:
//   DH = dh + B, DV = dv + B, S = s + B + p (a byte table, clamped at 0: a negative S never wins)
//   Z  = z + B + p = max3(DH_up, DV_left, S)                         ONE VIMNMX3.U16x2 per TWO cells
//   T  = Z - DH_up >= 0,  U = Z - DV_left >= 0                        borrow-free 32-bit subtractions
//   DV' = T + (B - p), DH' = U + (B - p);  Top iff T == 0, else Left iff U == 0, else Diagonal
// (get_direction's order Top > Left > Diagonal, enums.rs:18-28; global mode has no Beginning.)
// p = ext for every cell except (1, 1), which pays del (simple/mod.rs:72,88-92): an edge step.
//
// Geometry.  A long pair is latency-bound, not throughput-bound: the fill is a chain of anti-
// diagonal steps and the GPU is far from full, so the time is
//     n x (time of ONE step of ONE warp)  +  bands x (time from a band's first group in to its
//                                                     first group out)
// and nothing a band does later wins back what it lost at its start (every band runs at the same
// pace; measured with ALNR_DIFF_DEBUG time stamps).  The kernel therefore keeps both terms short:
// a lane owns TWO rows, packed in one 32-bit word on one anti-diagonal (row 2l at column tau, row
// 2l+1 at column tau - 1, tau = t - 2 l), so a step is one packed cell update (~25 instructions
// with everything around it) and a band is 64 rows whose last row starts after 63 steps, the
// minimum for 64 dependent rows.  Lane l + 1 receives lane l's last row by one shuffle per step.
// (A skew of three columns takes the shuffle's latency off the step's chain — 43.6 against 46.3 ns
// per step — but costs 31 more steps before every band's first hand-off; measured slower on every
// shape with more than a few bands.)  The band's last row goes to the next band through an L2-
// resident buffer of 16-bit values preset to a sentinel (data = flag, as in the tile pipeline),
// 8 columns per 16-byte store.  The ramp-up blocks of a band — on the critical path of every band
// below it — have their own instantiation without the column masks of the general edge code.
// Directions leave as the SAME blocked 4 x 4-tile words the tile pipeline writes: the two lanes of a
// 4-row strip exchange their step codes once per 8 steps and each un-shears one tile in registers,
// so every traceback is unchanged.
#pragma once
#include "kernels_tile.cuh"

namespace alnr {

constexpr uint32_t kDiffSentinel = 0xffffu;  // no DH value reaches it (DH <= 127 + 4 del)
constexpr uint32_t kDiffWake = 6;
constexpr uint32_t kDiffBandRows = 64;

struct DiffArgs {
  const uint8_t *q, *t;
  uint32_t n, m;
  uint32_t wtab[4];        // wtab[tsym] = 4 bytes S(tsym, qsym = 0..3) = clamp(s + B + p, 0, 127)
  int del, ext;
  uint32_t nbands;
  uint16_t *rows;          // [nbands][npad]: row b = DH of the row above band b, preset to 0xffff
                           // (row 0 is not used: the first band computes its boundary)
  uint32_t npad;           // n rounded up to a multiple of 64
  uint32_t *D;             // blocked 4 x 4 layout, or nullptr
  uint32_t dcols;
  int *corner_acc;         // sum over rows of dv[y][n] (zeroed by the host)
  unsigned long long *stamps;  // ALNR_DIFF_DEBUG: 8 time stamps per band, or nullptr
};
#ifdef ALNR_DIFF_DEBUG
#define DIFF_STAMP(k) do { if (lane == 0 && a.stamps) a.stamps[(size_t)band * 8 + (k)] = tile_now(); } while (0)
#else
#define DIFF_STAMP(k) do { } while (0)
#endif

// byte permute with sign replication (selector nibble bit 3): picks a score byte into a 16-bit half
__device__ __forceinline__ uint32_t diff_prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
  return d;
}

// 8 boundary values (columns col .. col + 7, col a multiple of 8), read from L2 by every lane.
// ld.relaxed.gpu, not ld.global.cg: ptxas may treat a weak load in a polling loop as loop-invariant
// (it did: the reload loop became a single reload once the loop body held nothing else).
__device__ __forceinline__ uint4 diff_ld_group(const uint16_t *row, uint32_t col) {
  uint4 g;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(g.x), "=r"(g.y), "=r"(g.z), "=r"(g.w)
               : "l"(row + col) : "memory");
  return g;
}

__device__ __forceinline__ uint32_t diff_ld_first(const uint16_t *row) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(row) : "memory");
  return v & 0xffffu;
}

// Per-lane state of a band sweep.
struct DiffLane {
  uint32_t DV, DH;      // the lane's rows (0, 1): DV at their previous column, DH of the previous step
  uint32_t Wt0, Wt1;    // the 4 scores of each row's target symbol, one byte per query symbol
  // Query symbols of the current 8-step block as READY-MADE selector bytes.  Row 0 sits at column
  // tau, row 1 at tau - 1: the step's score word is prmt(Wt0, Wt1, sel) with
  //   sel byte 0 = q[tau] * 0x11 + 0x80   (byte q of Wt0 into the low half, zero above it: the sign-
  //                                        replicating nibble 8|q reads 0 from a score 0..127)
  //   sel byte 1 = q[tau-1] * 0x11 + 0xC4 (byte q of Wt1 into the high half, zero above it)
  // A* holds the first form for columns tau0 .. tau0+7, B* the second for tau0-4 .. tau0+7, so sel is
  // one compile-time byte permute.  No memory access and no arithmetic on the step's chain.
  uint32_t A1, A2, B0, B1, B2;
  uint32_t hv[4];       // lane 31: 8 DH values of the band's last row being collected
};

__device__ __forceinline__ uint32_t diff_codes(uint32_t T, uint32_t U) {
  // per half: 0 Top (T == 0), 1 Left (T > 0, U == 0), 2 Diagonal  =  min(T, 1) + min(T, U, 1)
  const uint32_t one2 = 0x00010001u;
  return __vminu2(T, one2) + __vimin3_u16x2(T, U, one2);
}

// One step: the packed cell update of the lane's two rows.  t = band-local step, tau = t - 2 lane =
// column (0-based) of row 0, U = t & 7 (compile time).  upw: the word whose half (chosen by upsel)
// is the DH of the row above the lane at column tau.  Returns the two 2-bit codes (low half row 0,
// high half row 1).
// MODE 0: steady state.  MODE 1: ramp-up (t < 64, no column past n - 1, not cell (1,1)): a row that
// has not started computes garbage and its DV is (re)set when it starts — rstart = the step of the
// block at which the lane's row 0 starts (8 = not in this block), dv0w = its DV at column 0.
// MODE 2: general edge — rows outside [0, n) keep their DV, cell (1,1) pays del.
template <int U, int MODE>
__device__ __forceinline__ uint32_t diff_step(DiffLane &L, const DiffArgs &a, uint32_t t, uint32_t lane,
                                              bool first_band, uint32_t upw, uint32_t upsel, uint32_t K2,
                                              uint32_t rstart, uint32_t dv0w) {
  constexpr int i0 = 4 + U;  // index of q[tau0 + U] in the 12-byte stream tau0-4 .. tau0+7
  constexpr int wa = i0 >> 2, ba = i0 & 3, wb = (i0 - 1) >> 2, bb = (i0 - 1) & 3;
  const uint32_t Aw = wa == 1 ? L.A1 : L.A2;
  const uint32_t Bw = wb == 0 ? L.B0 : (wb == 1 ? L.B1 : L.B2);
  const uint32_t sel = __byte_perm(Aw, Bw, (uint32_t)ba | ((uint32_t)(4 + bb) << 4));
  const uint32_t S = diff_prmt(L.Wt0, L.Wt1, sel);
  const uint32_t up = diff_prmt(upw, L.DH, upsel);  // lo: row above the lane, hi: row 0 one step ago
  if (MODE == 1) {  // row 0 starts at an even step, row 1 one step later
    if ((U & 1) == 0 && rstart == (uint32_t)U) L.DV = dv0w;
    if ((U & 1) == 1 && rstart + 1u == (uint32_t)U) L.DV = __byte_perm(L.DV, dv0w, 0x7610);
  }
  uint32_t u = up, dv = L.DV, x11 = 0;
  if (MODE == 2 && first_band && lane == 0 && t == 0) {  // cell (1,1) pays del instead of ext (low half)
    x11 = (uint32_t)(a.del - a.ext);
    u -= x11; dv -= x11;
  }
  const uint32_t Z = __vimax3_u16x2(u, dv, S);
  const uint32_t T = Z - u, Uu = Z - dv;  // per half >= 0: no borrow between the halves
  uint32_t nDV = T + K2 - x11;
  if (MODE == 2) {  // rows outside [0, n) keep their DV (boundary value before, dv[y][n] after)
    const uint32_t tau = t - 2u * lane;  // wraps for lanes that have not started
    const uint32_t mk = ((tau < a.n) ? 0x0000ffffu : 0u) | ((tau - 1u < a.n) ? 0xffff0000u : 0u);
    nDV = (nDV & mk) | (L.DV & ~mk);
  }
  L.DV = nDV;
  L.DH = Uu + K2 - x11;
  return diff_codes(T, Uu);
}

// The 8 query symbols of columns col .. col + 7 as three aligned words and a funnel-shift count.
// Steady state: 0 <= col <= n (the sequence buffers carry 16 bytes of padding).
struct DiffQ8 {
  uint32_t lo, mid, hi, sh;
  uint32_t back;  // EDGE: the load was moved forward by this many columns (window straddles column 0)
};

template <bool EDGE>
__device__ __forceinline__ DiffQ8 diff_q8_issue(const uint8_t *q, uint32_t n, int col) {
  DiffQ8 r;
  int c = col;
  r.back = 0;
  if (EDGE) {  // columns outside [0, n) belong to masked cells: any bytes do, but never read outside
    c = col < 0 ? 0 : col;
    if ((uint32_t)c >= n) c = 0;
    r.back = (col < 0 && col > -8) ? (uint32_t)(-col) : 0u;
  }
  const uintptr_t adr = reinterpret_cast<uintptr_t>(q + c);
  const uint32_t *w = reinterpret_cast<const uint32_t *>(adr & ~(uintptr_t)3);
  r.lo = __ldg(w); r.mid = __ldg(w + 1); r.hi = __ldg(w + 2);
  r.sh = (uint32_t)(adr & 3) * 8u;
  return r;
}

// Turn the loaded words into the block's selector bytes (called when the block ENDS: the funnel
// shift would otherwise wait for the loads).
template <bool EDGE>
__device__ __forceinline__ void diff_q8_finish(DiffLane &L, const DiffQ8 &r) {
  uint32_t w1 = __funnelshift_r(r.lo, r.mid, r.sh), w2 = __funnelshift_r(r.mid, r.hi, r.sh);
  if (EDGE && r.back) {  // the bytes of column 0 onward start `back` bytes into the window
    const unsigned long long v = (((unsigned long long)w2 << 32) | w1) << (8u * r.back);
    w1 = (uint32_t)v; w2 = (uint32_t)(v >> 32);
  }
  L.B0 = L.B2;
  // padding bytes past the query's end may hold anything: the mask keeps them out of the nibbles
  L.A1 = (w1 & 0x03030303u) * 0x11u + 0x80808080u;
  L.A2 = (w2 & 0x03030303u) * 0x11u + 0x80808080u;
  L.B1 = L.A1 + 0x44444444u;
  L.B2 = L.A2 + 0x44444444u;
}

// What a band sweep carries from block to block besides the lane state.
struct DiffBand {
  uint32_t lane, K2, m0, m1, upsel_even;
  bool first_band, store_d, hand_out;
  uint32_t *dstrip;        // the 4-row strip of the lane pair in D, minus the lane's tile offset toff
  const uint16_t *row_in;  // DH of the row above the band
  uint16_t *row_out;       // DH of the band's last row, minus the 64 columns lane 31 is behind
  uint32_t shf;            // DH word of the lane above after the previous step
  uint4 bq0;               // boundary group of the next block (requested five steps ahead)
  uint32_t toff;           // un-shearing: tile column of the lane's tile = t0 / 4 - toff
  uint32_t Sprev;          // un-shearing: step bytes of the previous block's steps 4..7
  uint32_t hprev;          // DH after the previous block's last step (hand-off, two values per PRMT)
  uint32_t dv0w;           // the lane's DV at column 0 (ramp-up)
};

// One block of 8 steps.  t0 = band-local step of the block's first step (a multiple of 8).
// MODE as in diff_step; EDGE = MODE != 0 for everything around the steps.
template <int MODE>
__device__ __forceinline__ void diff_block(DiffLane &L, DiffBand &W, const DiffArgs &a, uint32_t t0, uint32_t n8) {
  constexpr bool EDGE = MODE != 0;
  const uint32_t lane = W.lane;
  // symbols of the NEXT block: issued now, turned into selector bytes when this block ends
  const DiffQ8 nq = diff_q8_issue<EDGE>(a.q, a.n, (int)(t0 + 8u) - 2 * (int)lane);
  // Boundary: lane 0 consumes the row above the band, 8 columns per block = one 16-byte group of the
  // producer (stored whole: the first and the last value are tested).  Every lane loads the same
  // group — one transaction, and the test is uniform.  The group was requested five steps ago.
  // The first band computes its boundary instead (a row written before the launch has left L2 by the
  // time it is read: ~850 cycles per group, measured 27 % of the band's time).
  uint4 g = W.bq0;
  if (W.first_band) {
    // row 0 of the matrix: dh[0][x] = -del for x < n and -2 del for x = n (H[0][n] = -(n+1) del,
    // simple/mod.rs:69); + B.  Index = 0-based column; a steady block can end with column n - 1.
    const uint32_t d1 = (uint32_t)a.del, d2 = d1 * 0x00010001u;
    g = make_uint4(d2, d2, d2, (t0 + 8u == a.n) ? d1 : d2);
    if (MODE == 2) {
      auto dh0 = [&](uint32_t col) { return (col + 1u == a.n) ? 0u : d1; };
      g = make_uint4(dh0(t0) | (dh0(t0 + 1) << 16), dh0(t0 + 2) | (dh0(t0 + 3) << 16),
                     dh0(t0 + 4) | (dh0(t0 + 5) << 16), dh0(t0 + 6) | (dh0(t0 + 7) << 16));
    }
  } else if (MODE != 2 || t0 < n8) {
    while ((g.x & 0xffffu) == kDiffSentinel || (g.w >> 16) == kDiffSentinel) g = diff_ld_group(W.row_in, t0);
  } else {
    g = make_uint4(0, 0, 0, 0);  // past the last column: masked cells, and never a sentinel in the arithmetic
  }
  // ramp-up: the step of this block at which the lane's row 0 starts (its column 0), 8 = none
  const uint32_t rstart = (MODE == 1 && t0 == ((2u * lane) & ~7u)) ? ((2u * lane) & 7u) : 8u;

  // codes: P collects two steps (step u at bits 8 (u & 1) of each half, odd lanes 4 bits higher),
  // X[j] = P | P >> 14 has the two step nibbles (row 0 | row 1 << 2) in its two low bytes
  uint32_t X[4], P = 0, dh_last = W.hprev;
#define ALNR_DIFF_STEP(U)                                                                               \
  {                                                                                                     \
    const uint32_t bw = (U) < 2 ? g.x : (U) < 4 ? g.y : (U) < 6 ? g.z : g.w;                            \
    const uint32_t upw = lane == 0 ? bw : W.shf;                                                        \
    const uint32_t c = diff_step<U, MODE>(L, a, t0 + (U), lane, W.first_band, upw,                      \
                                          ((U) & 1) ? 0x5432u : W.upsel_even, W.K2, rstart, W.dv0w);    \
    W.shf = __shfl_up_sync(0xffffffffu, L.DH, 1);                                                       \
    if (((U) & 1) == 0) {                                                                               \
      P = c * W.m0;                                                                                     \
      /* hand-off: the last row's DH (high half) of steps U - 1 and U: columns t - 64 and t - 63,  */  \
      /* positions U and U + 1 of their group of 8                                                 */  \
      L.hv[(U) >> 1] = __byte_perm(dh_last, L.DH, 0x7632);                                              \
    } else {                                                                                            \
      P += c * W.m1;                                                                                    \
      X[(U) >> 1] = P | (P >> 14);                                                                      \
    }                                                                                                   \
    dh_last = L.DH;                                                                                     \
    if ((U) == 6 && W.hand_out) {                                                                       \
      /* A group is half of a 32-byte sector.  A read of a sector that L2 holds only half of (the   */ \
      /* other half still the preset in DRAM) waits for DRAM: the first half is therefore stored   */ \
      /* together with a sentinel second half, so that L2 always owns whole sectors.               */ \
      /* row_out + t0 = the group's first column t0 - 64; wraps while ramping up.                  */ \
      /* The consumer asks for a group before it exists when it follows closely (every band's      */ \
      /* ramp-up does): the NEXT sector is therefore written as sentinels too, or that early read   */ \
      /* would fetch the preset from DRAM (~850 cycles; the rows outgrow L2 from n ~ 40 000 on).     */ \
      const uint4 s4 = make_uint4(~0u, ~0u, ~0u, ~0u);                                                  \
      if (MODE == 0) {                                                                                  \
        if (!(t0 & 8u)) {                                                                               \
          __stcg(reinterpret_cast<uint4 *>(W.row_out + t0 + 8u), s4);                                   \
          __stcg(reinterpret_cast<uint4 *>(W.row_out + t0 + 16u), s4);                                  \
          __stcg(reinterpret_cast<uint4 *>(W.row_out + t0 + 24u), s4);                                  \
        }                                                                                               \
        __stcg(reinterpret_cast<uint4 *>(W.row_out + t0), make_uint4(L.hv[0], L.hv[1], L.hv[2], L.hv[3])); \
      } else if (t0 - 64u < a.npad) { /* values of masked columns are garbage: never the sentinel */   \
        const uint32_t cap = 0xfffefffeu;                                                               \
        if (!(t0 & 8u)) {                                                                               \
          __stcg(reinterpret_cast<uint4 *>(W.row_out + t0 + 8u), s4);                                   \
          if (t0 - 64u + 32u <= a.npad) {                                                               \
            __stcg(reinterpret_cast<uint4 *>(W.row_out + t0 + 16u), s4);                                \
            __stcg(reinterpret_cast<uint4 *>(W.row_out + t0 + 24u), s4);                                \
          }                                                                                             \
        }                                                                                               \
        __stcg(reinterpret_cast<uint4 *>(W.row_out + t0),                                               \
               make_uint4(__vminu2(L.hv[0], cap), __vminu2(L.hv[1], cap), __vminu2(L.hv[2], cap),       \
                          __vminu2(L.hv[3], cap)));                                                     \
      }                                                                                                 \
    }                                                                                                   \
  }
  ALNR_DIFF_STEP(0) ALNR_DIFF_STEP(1) ALNR_DIFF_STEP(2)
  // The next block's group is requested HERE, not at the block's start: an L2 hit takes ~320 cycles,
  // about four steps, and a band that follows its producer closely would otherwise ask a block too
  // early, read the preset and pay a reload at every block of its ramp-up — which every band below
  // it inherits (the rows carry slack for the reads past their end).
  W.bq0 = diff_ld_group(W.row_in, t0 + 8u);
  ALNR_DIFF_STEP(3) ALNR_DIFF_STEP(4) ALNR_DIFF_STEP(5) ALNR_DIFF_STEP(6) ALNR_DIFF_STEP(7)
#undef ALNR_DIFF_STEP
  W.hprev = L.DH;

  // ---- directions.  Step bytes B(u) = even lane's nibble | odd lane's nibble << 4: row r of the
  // 4-row strip at bits 2r, at column tauE + u - r (tauE = the even lane's tau, a multiple of 4 at
  // the block's start).  Byte i of the blocked 4 x 4 tile word of columns c .. c+3 is therefore
  //   (B[c+i] & 0x03) | (B[c+i+1] & 0x0c) | (B[c+i+2] & 0x30) | (B[c+i+3] & 0xc0)
  // with B indexed by the column of row 0.  Per block the strip finishes two tiles: its even lane
  // builds columns tauE-4 .. tauE-1 (previous block's steps 4..7 + this block's 0..3), its odd lane
  // tauE .. tauE+3 (this block's steps 0..7).
  const uint32_t Alo = __byte_perm(X[0], X[1], 0x5410), Ahi = __byte_perm(X[2], X[3], 0x5410);
  const uint32_t Sa = Alo | __shfl_xor_sync(0xffffffffu, Alo, 1), Sb = Ahi | __shfl_xor_sync(0xffffffffu, Ahi, 1);
  const bool odd = (lane & 1u) != 0;
  const uint32_t S0 = odd ? Sa : W.Sprev, S1 = odd ? Sb : Sa;
  const uint32_t X1 = __byte_perm(S0, S1, 0x4321), X2 = __byte_perm(S0, S1, 0x5432), X3 = __byte_perm(S0, S1, 0x6543);
  uint32_t w = (S0 & 0x03030303u) | (X1 & 0x0c0c0c0cu);
  w = w | (X2 & 0x30303030u);
  w = w | (X3 & 0xc0c0c0c0u);
  const uint32_t tix = (t0 >> 2) - W.toff;  // tile column; wraps while the lanes ramp up
  if (W.store_d && (!EDGE || tix < (a.n + 3u) / 4u)) __stcs(W.dstrip + (t0 >> 2), w);
  W.Sprev = Sb;
  diff_q8_finish<EDGE>(L, nq);
}

// One persistent launch: band = blockIdx.x, blockIdx.x + gridDim.x, ...; a waiter only depends on
// lower-numbered bands.  One warp per block; shared memory: 4 words (the S tables).
__global__ void __launch_bounds__(32) fill_diff_kernel(DiffArgs a) {
  __shared__ uint32_t swt[4];
  const uint32_t lane = threadIdx.x;
  if (lane < 4) swt[lane] = a.wtab[lane];
  __syncwarp();
  const uint32_t n = a.n, m = a.m;
  const uint32_t B = 2u * (uint32_t)a.del;
  const uint32_t n8 = (n + 7u) / 8u * 8u;  // the producer stores whole groups of 8 columns
  const uint32_t steps = n8 + 72u;         // last tile of lane 31 and last hand-off group included
  int corner_part = 0;

  for (uint32_t band = blockIdx.x; band < a.nbands; band += gridDim.x) {
    const uint32_t y0 = band * kDiffBandRows + lane * 2u;  // 0-based index of the lane's first row
    const uint32_t j = lane >> 1;                          // the lane pair's 4-row strip of the band
    DiffLane L;
    DiffBand W;
    W.lane = lane;
    W.K2 = (B - (uint32_t)a.ext) * 0x00010001u;
    W.m0 = 1u << (4u * (lane & 1u));
    W.m1 = 256u << (4u * (lane & 1u));
    asm volatile("" : "+r"(W.m0), "+r"(W.m1));  // opaque: a multiply-add on the fma pipe, not a shift on the alu pipe
    W.upsel_even = lane == 0 ? 0x5410u : 0x5432u;  // lane 0 reads a boundary word: even column = low half
    W.first_band = band == 0;
    W.store_d = a.D != nullptr && band * kDiffBandRows + j * 4u < m;
    W.hand_out = lane == 31 && band + 1 < a.nbands;
    W.row_in = a.rows + (size_t)band * a.npad;
    W.row_out = a.rows + (size_t)(band + 1) * a.npad - 64;  // only dereferenced when hand_out
    W.shf = 0; W.Sprev = 0; W.hprev = 0;
    W.toff = j + 1u - (lane & 1u);
    W.dstrip = a.D + (size_t)(band * (kDiffBandRows / 4u) + j) * a.dcols - W.toff;
    // opaque: kept in registers instead of being re-derived from the kernel parameters in every block
    asm volatile("" : "+l"(W.dstrip), "+l"(W.row_out), "+l"(W.row_in));
    L.Wt0 = swt[(y0 < m) ? (uint32_t)a.t[y0] & 3u : 0u];
    L.Wt1 = swt[(y0 + 1 < m) ? (uint32_t)a.t[y0 + 1] & 3u : 0u];
    // column 0: dv[y][0] = -del, and -2 del on row m (H[m][0] = -(m+1) del, simple/mod.rs:70); + B
    auto dv0 = [&](uint32_t r) { return (y0 + r + 1 == m) ? 0u : (uint32_t)a.del; };
    W.dv0w = dv0(0) | (dv0(1) << 16);
    L.DV = W.dv0w;
    L.DH = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) L.hv[k] = 0;
    // selector bytes of the first block: columns tau0 - 4 .. tau0 + 7 with tau0 = -2 lane
    L.B2 = 0xC4C4C4C4u;  // becomes B0: columns before the matrix
    diff_q8_finish<true>(L, diff_q8_issue<true>(a.q, n, -2 * (int)lane));

    // a waiting band must not take issue slots from working ones: sleep until band b - kDiffWake
    // has handed on its first group, then poll the own boundary row
    DIFF_STAMP(0);
    if (band > kDiffWake) {
      const uint16_t *far = a.rows + (size_t)(band - kDiffWake + 1) * a.npad;
      while (diff_ld_first(far) == kDiffSentinel) __nanosleep(500);
    }
    DIFF_STAMP(1);
    W.bq0 = diff_ld_group(W.row_in, 0);
    for (uint32_t t0 = 0; t0 < steps; t0 += 8u) {
#ifdef ALNR_DIFF_DEBUG
      if (t0 == 8u) DIFF_STAMP(2);
      if (t0 == 64u) DIFF_STAMP(3);
      if (t0 == 72u) DIFF_STAMP(4);
      if (t0 == 8192u) DIFF_STAMP(5);
#endif
      // steady state: every row of every lane is inside the matrix for all 8 steps (lane 31's row 1
      // is at column t - 63), every tile and hand-off group is in range.  Ramp-up: no column past
      // n - 1 yet; the block with cell (1,1) takes the general edge code.
      if (t0 + 8u <= n && t0 >= 64u) diff_block<0>(L, W, a, t0, n8);
      else if (t0 + 8u <= n && (t0 > 0u || band > 0u)) diff_block<1>(L, W, a, t0, n8);
      else diff_block<2>(L, W, a, t0, n8);
    }
    DIFF_STAMP(6);
    // H[m][n] = H[0][n] + sum over rows of dv[y][n]: every row's DV now holds dv[y][n] + B
    if (y0 < m) corner_part += (int)(L.DV & 0xffffu) - (int)B;
    if (y0 + 1 < m) corner_part += (int)(L.DV >> 16) - (int)B;
    __syncwarp();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) corner_part += __shfl_xor_sync(0xffffffffu, corner_part, o);
  if (lane == 0 && corner_part) atomicAdd(a.corner_acc, corner_part);
}

// corner = H[0][n] + sum of dv = -(n+1) del + acc; f = 0, the traceback starts at (m, n)
static __global__ void diff_finish_kernel(DiffArgs a, FillOut fo, uint32_t pair) {
  fo.corner[pair] = (double)(-(int)(a.n + 1) * a.del + *a.corner_acc);
  fo.f[pair] = 0.0;  // simple/mod.rs:139
  fo.start[2 * (size_t)pair] = a.m;
  fo.start[2 * (size_t)pair + 1] = a.n;
}

}  // namespace alnr
