// kernels_exact.cuh — the EXACT column-major fill: one pair per thread, iterating cells in
// the reference's own order (x outer / y inner, aligner-core/src/simple/mod.rs:74-97 and
// :187-210), so the loop-carried `penalty` state (simple/mod.rs:72,88-92 / :185,201-205)
// needs no speculation.  It is the universal path: any lengths, any del/ext, integer or
// FP64 (heuristic matrices are non-integral => f64).  The previous score column lives in
// an HBM/L2 scratch laid out [y][lane] so a warp's accesses coalesce; directions go to the
// 2-bit D scratch (common.cuh).  The fast s16x2 strip kernel (kernels_strip.cuh) handles
// the short-read batches; pairs whose speculation fails there are re-run here.
#pragma once
#include <cfloat>
#include "common.cuh"

namespace alnr {

template <typename T> struct Num;
template <> struct Num<int> {
  // integer domain: exact because every reference value is an integer < 2^53 in f64.
  static __device__ __forceinline__ int max3(int a, int b, int c) { return max(max(a, b), c); }
  static __device__ __forceinline__ bool same(int mx, int c) { return mx == c; }  // |mx-c| < EPS
  static __device__ __forceinline__ bool is_zero(int v) { return v == 0; }
};
template <> struct Num<double> {
  // enums.rs:19: f64::max(f64::max(top,left),diagonal); fmax ignores NaN like f64::max.
  static __device__ __forceinline__ double max3(double a, double b, double c) {
    return fmax(fmax(a, b), c);
  }
  // enums.rs:21,23: (max - cand).abs() < f64::EPSILON  (__dsub_rn: never contracted)
  static __device__ __forceinline__ bool same(double mx, double c) {
    return fabs(__dsub_rn(mx, c)) < DBL_EPSILON;
  }
  static __device__ __forceinline__ bool is_zero(double v) { return v == 0.0; }  // enums.rs:37
};
__device__ __forceinline__ int nsub(int a, int b) { return a - b; }
__device__ __forceinline__ int nadd(int a, int b) { return a + b; }
__device__ __forceinline__ double nsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double nadd(double a, double b) { return __dadd_rn(a, b); }

template <typename T>
struct ExactArgs {
  ChunkView cv;
  int mode;          // 0 global, 1 local
  int V;
  T del, ext;
  const T *matrix;   // device, V*V, [t][q]
  T *hcol;           // scratch [(group*(hrows) + y)*32 + lane], hrows = m_max + 1
  uint32_t hrows;
  uint32_t *D;
  DLayout dl;
  FillOut fo;
  double *H_out;     // optional, single-pair calls only: (m+1)*(n+1) row-major
  uint8_t *D_out;    // optional, 1 byte per cell
  const uint32_t *list = nullptr;        // optional: chunk slots to process (strip leftovers)
  const uint32_t *list_count = nullptr;
  const T *item_matrices = nullptr;      // optional: one V*V matrix per result index (heuristic search)
};

// grid: ceil(count/128) blocks of 128 threads; smem: V*V*sizeof(T)
template <typename T>
__global__ void __launch_bounds__(128) fill_exact_kernel(ExactArgs<T> a) {
  extern __shared__ unsigned char smem_raw[];
  T *smat_s = reinterpret_cast<T *>(smem_raw);
  // PWM view (common.cuh): the matrix is the 4 x L profile, indexed [target symbol][column]
  const bool pwm = a.cv.pwm_len != 0;
  const int stride = pwm ? (int)a.cv.pwm_len : a.V, elems = pwm ? 4 * stride : a.V * a.V;
  if (!a.item_matrices) for (int i = threadIdx.x; i < elems; i += blockDim.x) smat_s[i] = a.matrix[i];
  __syncthreads();

  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t slot = idx;
  if (a.list) {
    if (idx >= *a.list_count) return;
    slot = a.list[idx];
  }
  if (slot >= a.cv.count) return;
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first + slot] : a.cv.first + slot;
  // rejected by validation (empty / bad symbol); 11 may be left over from a speculative pass
  if (a.fo.status[pair] != 0 && !(a.list && a.fo.status[pair] == 11)) return;
  if (a.list) a.fo.status[pair] = 0;

  const uint32_t sp = a.cv.seq(pair);
  const uint8_t *q, *t;
  uint32_t n, m;
  a.cv.pair_dims(sp, q, t, n, m);
  const T *smat = a.item_matrices ? a.item_matrices + (size_t)pair * elems : smat_s;
  const bool glob = a.mode == 0;
  const uint32_t group = idx >> 5, lane = idx & 31;  // score-column scratch is per thread
  T *hc = a.hcol + (uint64_t)group * a.hrows * 32 + lane;  // hc[y*32]
  const uint32_t rpw = a.dl.rows_per_word;
  const size_t W = (size_t)n + 1;

  // column 0 (simple/mod.rs:64-67, :70 global; zeros in local)
  for (uint32_t y = 0; y <= m; ++y) {
    T v = T(0);
    if (glob && y > 0) v = (y == m) ? -(T(m) + T(1)) * a.del : -T(y) * a.del;
    hc[(size_t)y * 32] = v;
    if (a.H_out) a.H_out[y * W] = (double)v;
    if (a.D_out) a.D_out[y * W] = glob ? (y == 0 ? DIR_BEGINNING : DIR_TOP) : DIR_BEGINNING;
  }

  T pen = a.del;
  T best = T(0);
  uint32_t by = 0, bx = 0;
  T last = T(0);
  for (uint32_t x = 1; x <= n; ++x) {
    const int qs = pwm ? (int)(x - 1) : q[x - 1];  // pwm/mod.rs:62
    // row 0 of this column (simple/mod.rs:59-62, :69)
    T hup = T(0);
    if (glob) hup = (x == n) ? -(T(n) + T(1)) * a.del : -T(x) * a.del;
    T diag = hc[0];
    hc[0] = hup;
    if (a.H_out) a.H_out[x] = (double)hup;
    if (a.D_out) a.D_out[x] = glob ? DIR_LEFT : DIR_BEGINNING;
    uint32_t dword = 0;
    for (uint32_t y = 1; y <= m; ++y) {
      const T left = hc[(size_t)y * 32];
      const T s = smat[(int)t[y - 1] * stride + qs];
      const T top = nsub(hup, pen);
      const T lft = nsub(left, pen);
      const T dg = nadd(diag, s);
      const T mx = Num<T>::max3(top, lft, dg);
      int d;
      if (!glob && Num<T>::is_zero(mx)) d = DIR_BEGINNING;       // enums.rs:37
      else if (Num<T>::same(mx, top)) d = DIR_TOP;               // enums.rs:21/:39
      else if (Num<T>::same(mx, lft)) d = DIR_LEFT;              // enums.rs:23/:41
      else d = DIR_DIAGONAL;
      pen = (d != DIR_BEGINNING) ? a.ext : a.del;                // simple/mod.rs:88-92
      diag = left;
      hup = mx;
      hc[(size_t)y * 32] = mx;
      const uint32_t r = (y - 1) % rpw;
      dword |= a.dl.put((uint32_t)d, r);
      if (r == rpw - 1 || y == m) {
        if (a.D) a.D[a.dl.word(slot, (y - 1) / rpw, x - 1)] = dword;  // nullptr: score-only call
        dword = 0;
      }
      if (a.H_out) a.H_out[y * W + x] = (double)mx;
      if (a.D_out) a.D_out[y * W + x] = (uint8_t)d;
      // ndarray-stats argmax: first maximum in row-major order (smallest y, then x)
      if (!glob && (mx > best || (mx == best && y < by))) { best = mx; by = y; bx = x; }
    }
    last = hup;
  }

  a.fo.corner[pair] = (double)last;
  if (glob) {
    a.fo.f[pair] = 0.0;  // simple/mod.rs:139
    a.fo.start[2 * (size_t)pair] = m;
    a.fo.start[2 * (size_t)pair + 1] = n;
  } else {
    a.fo.f[pair] = (double)best;  // simple/mod.rs:247
    a.fo.start[2 * (size_t)pair] = by;
    a.fo.start[2 * (size_t)pair + 1] = bx;
    // ALNR_NO_POSITIVE_CELL (simple/mod.rs:214); the PWM aligner returns an empty alignment
    if ((by == 0 || bx == 0) && !pwm) a.fo.status[pair] = 11;
  }
}

}  // namespace alnr
