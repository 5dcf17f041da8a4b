// oracle_matrix.cpp — TEST INFRASTRUCTURE (see oracle.h).
// Restatement of aligner-helpers/src/matrices/mod.rs:8-80 (transform_matrix,
// get_threshold, cross_product) plus the three pieces of THIRD-PARTY arithmetic it
// relies on, whose sources are NOT under /root/reference and not on this machine:
//   * ndarray 0.15.4 (Cargo.lock:1266-1267)  ArrayBase::sum -> numeric_util::unrolled_fold
//   * ndarray-stats 0.5.0 (Cargo.lock:1291-1292)  DeviationExt::l2_dist
//   * roots 0.0.7 (Cargo.lock:1781-1782)  find_roots_quadratic
// They are restated from their published algorithms.  No reference test pins them:
// PARITY UNPINNED for transform_matrix bit patterns.
#include "oracle.h"

#include <cmath>
#include <vector>

// ndarray 0.15 numeric_util::unrolled_fold(xs, zero, add): eight running partial sums
// over chunks of 8, combined as ((((0+(p0+p4))+(p1+p5))+(p2+p6))+(p3+p7)), then the
// (<8) tail elements added one by one.
extern "C" double orc_ndarray_sum(const double *x, size_t len) {
  double p0 = 0, p1 = 0, p2 = 0, p3 = 0, p4 = 0, p5 = 0, p6 = 0, p7 = 0;
  size_t i = 0;
  for (; i + 8 <= len; i += 8) {
    p0 = p0 + x[i + 0]; p1 = p1 + x[i + 1]; p2 = p2 + x[i + 2]; p3 = p3 + x[i + 3];
    p4 = p4 + x[i + 4]; p5 = p5 + x[i + 5]; p6 = p6 + x[i + 6]; p7 = p7 + x[i + 7];
  }
  double acc = 0.0;
  acc = acc + (p0 + p4);
  acc = acc + (p1 + p5);
  acc = acc + (p2 + p6);
  acc = acc + (p3 + p7);
  for (; i < len; ++i) acc = acc + x[i];
  return acc;
}

// ndarray-stats 0.5 sq_l2_dist: sequential  sum += (a-b)*(a-b)  in memory order; sqrt.
extern "C" double orc_l2_dist(const double *a, const double *b, size_t len) {
  double s = 0.0;
  for (size_t i = 0; i < len; ++i) {
    double d = a[i] - b[i];
    s += d * d;
  }
  return std::sqrt(s);
}

// roots 0.0.7 find_roots_quadratic(a2, a1, a0)
extern "C" int orc_find_roots_quadratic(double a2, double a1, double a0, double *roots) {
  if (a2 == 0.0) {  // linear fallback: a1*x + a0 = 0
    if (a1 == 0.0) return 0;
    roots[0] = -a0 / a1;
    return 1;
  }
  double disc = a1 * a1 - 4.0 * a2 * a0;
  if (disc < 0.0) return 0;
  double a2x2 = 2.0 * a2;
  if (disc == 0.0) {
    roots[0] = -a1 / a2x2;
    return 1;
  }
  double sq = std::sqrt(disc);
  double same_sign, diff_sign;
  if (a1 < 0.0) { same_sign = -a1 + sq; diff_sign = -a1 - sq; }
  else          { same_sign = -a1 - sq; diff_sign = -a1 + sq; }
  double x1, x2;
  if (std::fabs(same_sign) > std::fabs(a2x2)) {
    double a0x2 = 2.0 * a0;
    if (std::fabs(diff_sign) > std::fabs(a2x2)) { x1 = a0x2 / same_sign; x2 = a0x2 / diff_sign; }
    else { x1 = a0x2 / same_sign; x2 = same_sign / a2x2; }
  } else {
    x1 = diff_sign / a2x2; x2 = same_sign / a2x2;
  }
  if (x1 < x2) { roots[0] = x1; roots[1] = x2; }
  else         { roots[0] = x2; roots[1] = x1; }
  return 2;
}

extern "C" double orc_get_threshold(int dim) {  // matrices/mod.rs:8-17
  switch (dim) {
    case 20: return 22.6;
    case 21: return 23.1;
    case 22: return 23.6;
    case 23: return 24.1;
    case 24: return 24.6;
    default: return 0.0;
  }
}

extern "C" int orc_transform_matrix(const double *M, int R, int C, double k_d,
                                    double r_squared, const double *freq, double *out) {
  const size_t N = (size_t)R * (size_t)C;
  std::vector<double> p(N), tmp(N), e(N);
  const double fcol = 1.0 / (double)C;  // mod.rs:25-27
  for (int y = 0; y < R; ++y)           // mod.rs:70-80 cross_product
    for (int x = 0; x < C; ++x) p[(size_t)y * C + x] = freq[y] * fcol;

  for (size_t i = 0; i < N; ++i) tmp[i] = p[i] * p[i];
  const double p_squared = orc_ndarray_sum(tmp.data(), N);  // :31
  for (size_t i = 0; i < N; ++i) tmp[i] = p[i] * M[i];
  const double k_0 = orc_ndarray_sum(tmp.data(), N);        // :33

  const double a = (k_d - k_0) / p_squared;  // :35
  const double b = k_d / p_squared;          // :36
  const double difference = a - b;           // :37

  for (size_t i = 0; i < N; ++i) e[i] = M[i] + p[i] * difference;
  for (size_t i = 0; i < N; ++i) tmp[i] = e[i] * e[i];
  const double denominator = orc_ndarray_sum(tmp.data(), N);  // :39
  for (size_t i = 0; i < N; ++i) tmp[i] = p[i] * e[i];
  const double a_coeff = (2.0 * b * orc_ndarray_sum(tmp.data(), N)) / denominator;  // :41
  const double b_coeff = (b * b * p_squared - r_squared) / denominator;              // :42

  double roots[2];
  int nr = orc_find_roots_quadratic(1.0, a_coeff, b_coeff, roots);  // :44
  if (nr == 0) return ORC_WRONG_MATRIX;                              // :45
  auto build = [&](double root, double *dst) {  // &p * b + root * (matrix + &p * (a - b))
    for (size_t i = 0; i < N; ++i) dst[i] = p[i] * b + root * (M[i] + p[i] * (a - b));
  };
  if (nr == 1) { build(roots[0], out); return ORC_OK; }               // :46
  if (roots[0] > 0.0 && roots[1] < 0.0) { build(roots[0], out); return ORC_OK; }  // :48-49
  if (roots[0] < 0.0 && roots[1] > 0.0) { build(roots[1], out); return ORC_OK; }  // :50-51
  std::vector<double> m1(N), m2(N);  // :53-63
  build(roots[0], m1.data());
  build(roots[1], m2.data());
  double d1 = orc_l2_dist(M, m1.data(), N);
  double d2 = orc_l2_dist(M, m2.data(), N);
  const std::vector<double> &pick = (d1 < d2) ? m1 : m2;
  for (size_t i = 0; i < N; ++i) out[i] = pick[i];
  return ORC_OK;
}
