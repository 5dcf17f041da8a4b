// transform.cuh — transform_matrix (aligner-helpers/src/matrices/mod.rs:19-68) as a
// __host__ __device__ routine, one matrix per thread, streaming (no temporaries).
//
// Every element-wise product/sum is rounded separately (the library is built with
// -fmad=false; host code with -ffp-contract=off) and the reductions follow the order of the
// crates the reference calls — ndarray 0.15.4 `sum()` (eight running partial sums over chunks of
// 8, combined ((((0+(p0+p4))+(p1+p5))+(p2+p6))+(p3+p7)), then the tail one by one) and
// ndarray-stats 0.5.0 `l2_dist` (sequential) — and roots 0.0.7 `find_roots_quadratic`.  Those
// crate sources are not in the reference tree: restated from their published algorithms,
// PARITY UNPINNED (see oracle/oracle_matrix.cpp, which this must match bit for bit).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace alnr {

// ndarray unrolled_fold of f(i) for i in [0, N)
template <class F>
__host__ __device__ inline double nd_sum(int N, F f) {
  double p0 = 0, p1 = 0, p2 = 0, p3 = 0, p4 = 0, p5 = 0, p6 = 0, p7 = 0;
  int i = 0;
  for (; i + 8 <= N; i += 8) {
    p0 = p0 + f(i + 0); p1 = p1 + f(i + 1); p2 = p2 + f(i + 2); p3 = p3 + f(i + 3);
    p4 = p4 + f(i + 4); p5 = p5 + f(i + 5); p6 = p6 + f(i + 6); p7 = p7 + f(i + 7);
  }
  double acc = 0.0;
  acc = acc + (p0 + p4);
  acc = acc + (p1 + p5);
  acc = acc + (p2 + p6);
  acc = acc + (p3 + p7);
  for (; i < N; ++i) acc = acc + f(i);
  return acc;
}

// roots 0.0.7 find_roots_quadratic(1, a1, a0); returns the number of roots, ascending.
__host__ __device__ inline int roots_quadratic_monic(double a1, double a0, double *r) {
  const double a2 = 1.0;
  const double disc = a1 * a1 - 4.0 * a2 * a0;
  if (disc < 0.0) return 0;
  const double a2x2 = 2.0 * a2;
  if (disc == 0.0) { r[0] = -a1 / a2x2; return 1; }
  const double sq = sqrt(disc);
  double same_sign, diff_sign;
  if (a1 < 0.0) { same_sign = -a1 + sq; diff_sign = -a1 - sq; }
  else          { same_sign = -a1 - sq; diff_sign = -a1 + sq; }
  double x1, x2;
  if (fabs(same_sign) > fabs(a2x2)) {
    const double a0x2 = 2.0 * a0;
    if (fabs(diff_sign) > fabs(a2x2)) { x1 = a0x2 / same_sign; x2 = a0x2 / diff_sign; }
    else { x1 = a0x2 / same_sign; x2 = same_sign / a2x2; }
  } else {
    x1 = diff_sign / a2x2; x2 = same_sign / a2x2;
  }
  if (x1 < x2) { r[0] = x1; r[1] = x2; } else { r[0] = x2; r[1] = x1; }
  return 2;
}

// out may not alias M.  Returns 0, or 9 (ALNR_WRONG_MATRIX) on Roots::No (mod.rs:45).
__host__ __device__ inline int transform_matrix_rc(const double *M, int R, int C, double k_d,
                                                   double r_squared, const double *freq, double *out) {
  const int N = R * C;
  const double fcol = 1.0 / (double)C;                                   // mod.rs:25-27
  auto p = [&](int i) { return freq[i / C] * fcol; };                    // cross_product, :70-80
  const double p_squared = nd_sum(N, [&](int i) { double v = p(i); return v * v; });  // :31
  const double k_0 = nd_sum(N, [&](int i) { return p(i) * M[i]; });      // :33
  const double a = (k_d - k_0) / p_squared;                              // :35
  const double b = k_d / p_squared;                                      // :36
  const double difference = a - b;                                       // :37
  auto e = [&](int i) { return M[i] + p(i) * difference; };
  const double denominator = nd_sum(N, [&](int i) { double v = e(i); return v * v; });   // :39
  const double a_coeff = (2.0 * b * nd_sum(N, [&](int i) { return p(i) * e(i); })) / denominator;  // :41
  const double b_coeff = (b * b * p_squared - r_squared) / denominator;  // :42
  double roots[2];
  const int nr = roots_quadratic_monic(a_coeff, b_coeff, roots);         // :44
  if (nr == 0) return 9;
  auto elem = [&](double root, int i) { return p(i) * b + root * (M[i] + p(i) * (a - b)); };
  double pick;
  if (nr == 1) pick = roots[0];                                          // :46
  else if (roots[0] > 0.0 && roots[1] < 0.0) pick = roots[0];            // :48-49
  else if (roots[0] < 0.0 && roots[1] > 0.0) pick = roots[1];            // :50-51
  else {                                                                 // :53-63
    double s1 = 0.0, s2 = 0.0;
    for (int i = 0; i < N; ++i) {
      const double d1 = M[i] - elem(roots[0], i), d2 = M[i] - elem(roots[1], i);
      s1 += d1 * d1;
      s2 += d2 * d2;
    }
    pick = (sqrt(s1) < sqrt(s2)) ? roots[0] : roots[1];
  }
  for (int i = 0; i < N; ++i) out[i] = elem(pick, i);
  return 0;
}

#ifdef __CUDACC__
// The same routine run by a GROUP of 8 consecutive lanes (gl = lane & 7, gmask = the group's lanes):
// ndarray's eight running partial sums are independent chains, so lane j keeps p_j; the combine
// and the tail follow the same order as nd_sum, every lane ends up with the same bits.  Element-
// wise work is strided over the group.  All 8 lanes must call it together.
template <class F>
__device__ inline double nd_sum_g8(int N, F f, unsigned gmask, int gl) {
  double p = 0.0;
  int i = 0;
  for (; i + 8 <= N; i += 8) p = p + f(i + gl);
  const double s = p + __shfl_down_sync(gmask, p, 4, 8);  // lanes 0..3: p_j + p_{j+4}
  double acc = 0.0;
#pragma unroll
  for (int j = 0; j < 4; ++j) acc = acc + __shfl_sync(gmask, s, j, 8);
  for (; i < N; ++i) acc = acc + f(i);
  return acc;
}

__device__ inline int transform_matrix_g8(const double *M, int R, int C, double k_d, double r_squared,
                                          const double *freq, double *out, unsigned gmask, int gl) {
  const int N = R * C;
  const double fcol = 1.0 / (double)C;
  auto p = [&](int i) { return freq[i / C] * fcol; };
  const double p_squared = nd_sum_g8(N, [&](int i) { double v = p(i); return v * v; }, gmask, gl);
  const double k_0 = nd_sum_g8(N, [&](int i) { return p(i) * M[i]; }, gmask, gl);
  const double a = (k_d - k_0) / p_squared;
  const double b = k_d / p_squared;
  const double difference = a - b;
  auto e = [&](int i) { return M[i] + p(i) * difference; };
  const double denominator = nd_sum_g8(N, [&](int i) { double v = e(i); return v * v; }, gmask, gl);
  const double a_coeff = (2.0 * b * nd_sum_g8(N, [&](int i) { return p(i) * e(i); }, gmask, gl)) / denominator;
  const double b_coeff = (b * b * p_squared - r_squared) / denominator;
  double roots[2];
  const int nr = roots_quadratic_monic(a_coeff, b_coeff, roots);
  if (nr == 0) return 9;
  auto elem = [&](double root, int i) { return p(i) * b + root * (M[i] + p(i) * (a - b)); };
  double pick;
  if (nr == 1) pick = roots[0];
  else if (roots[0] > 0.0 && roots[1] < 0.0) pick = roots[0];
  else if (roots[0] < 0.0 && roots[1] > 0.0) pick = roots[1];
  else {  // l2_dist is a sequential fold: every lane runs it (same bits), rare branch
    double s1 = 0.0, s2 = 0.0;
    for (int i = 0; i < N; ++i) {
      const double d1 = M[i] - elem(roots[0], i), d2 = M[i] - elem(roots[1], i);
      s1 += d1 * d1;
      s2 += d2 * d2;
    }
    pick = (sqrt(s1) < sqrt(s2)) ? roots[0] : roots[1];
  }
  for (int i = gl; i < N; i += 8) out[i] = elem(pick, i);
  return 0;
}
#endif

}  // namespace alnr
