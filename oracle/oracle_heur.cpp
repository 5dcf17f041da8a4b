// oracle_heur.cpp — TEST INFRASTRUCTURE (see oracle.h).  Restatement of
// aligner-core/src/alignment.rs:13-42 (frequency matrix, consensus line),
// aligner-core/src/heuristic/mod.rs:36-78 (HeuristicAligner::perform_alignment) and
// thread-pooled batch drivers used as parity checker and timed CPU baseline.
#include "oracle.h"

#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

extern "C" void orc_frequency_matrix(const uint8_t *qa, const uint8_t *ta, size_t len, int V,
                                     double *F) {
  for (int i = 0; i < V * V; ++i) F[i] = 0.0;
  for (size_t i = 0; i < len; ++i)  // alignment.rs:16-20
    if (qa[i] != ORC_BLANK && ta[i] != ORC_BLANK) F[(size_t)ta[i] * V + qa[i]] += 1.0;
}

extern "C" void orc_get_alignment(const uint8_t *qa, const uint8_t *ta, size_t len,
                                  const double *matrix, int V, uint8_t *out) {
  for (size_t i = 0; i < len; ++i) {  // alignment.rs:28-39
    if (qa[i] == ta[i]) out[i] = qa[i];
    else if (qa[i] != ORC_BLANK && ta[i] != ORC_BLANK &&
             matrix[(size_t)ta[i] * V + qa[i]] >= 0.0) out[i] = ORC_POS;
    else out[i] = ORC_BLANK;
  }
}

extern "C" int orc_heuristic(const uint8_t *q, size_t n, const uint8_t *t, size_t m,
                             double del, double ext, const double *matrix, int V, double kd,
                             double r_squared, const double *freq, uint8_t *q_aln,
                             uint8_t *t_aln, size_t aln_cap, size_t *aln_len,
                             uint64_t *coords, double *f, double *out_matrix, int *passes) {
  const size_t VV = (size_t)V * V;
  if (std::fabs(r_squared - 0.0) < DBL_EPSILON) r_squared = (double)VV;  // heuristic/mod.rs:48-50
  std::vector<double> T(VV), F(VV), T2(VV);
  int st = orc_transform_matrix(matrix, V, V, kd, r_squared, freq, T.data());  // :52-53
  if (st) return st;
  std::vector<uint8_t> qa(n + m + 2), ta(n + m + 2);
  double max_f = 0.0;
  int np = 0;
  for (;;) {  // :58-77
    size_t L;
    double fv;
    uint64_t c[4];
    st = orc_align_pair(1, q, n, t, m, del, ext, T.data(), V, nullptr, nullptr, qa.data(),
                        ta.data(), qa.size(), &L, c, &fv, nullptr);
    if (st) return st;
    ++np;
    if (fv > max_f) {  // :64
      max_f = fv;
      orc_frequency_matrix(qa.data(), ta.data(), L, V, F.data());
      st = orc_transform_matrix(F.data(), V, V, kd, r_squared, freq, T2.data());  // :66-72
      if (st) return st;
      T.swap(T2);
    } else {  // :73-76
      if (aln_len) *aln_len = L;
      if (q_aln || t_aln) {
        if (L > aln_cap) return ORC_BUFFER_TOO_SMALL;
        if (q_aln) std::memcpy(q_aln, qa.data(), L);
        if (t_aln) std::memcpy(t_aln, ta.data(), L);
      }
      if (coords) std::memcpy(coords, c, sizeof(c));
      if (f) *f = fv;
      if (out_matrix) std::memcpy(out_matrix, T.data(), VV * sizeof(double));
      if (passes) *passes = np;
      return ORC_OK;
    }
  }
}

namespace {
template <class Fn>
void parallel_for(size_t count, int threads, Fn fn) {
  if (threads < 1) threads = 1;
  std::atomic<size_t> next{0};
  auto worker = [&]() {
    for (;;) {
      size_t i0 = next.fetch_add(64);
      if (i0 >= count) break;
      size_t i1 = i0 + 64 < count ? i0 + 64 : count;
      for (size_t i = i0; i < i1; ++i) fn(i);
    }
  };
  std::vector<std::thread> pool;
  for (int k = 1; k < threads; ++k) pool.emplace_back(worker);
  worker();
  for (auto &th : pool) th.join();
}
}  // namespace

extern "C" int orc_align_batch(int mode, const uint8_t *qs, const uint64_t *qoff,
                               const uint8_t *ts, const uint64_t *toff, size_t count,
                               double del, double ext, const double *matrix, int V,
                               int threads, double *f, double *corner, uint64_t *coords,
                               uint64_t *aln_len, uint8_t *q_aln_all, uint8_t *t_aln_all,
                               size_t slot, int *status) {
  std::atomic<int> worst{0};
  parallel_for(count, threads, [&](size_t i) {
    size_t n = qoff[i + 1] - qoff[i], m = toff[i + 1] - toff[i];
    size_t L = 0;
    double fv = 0, cv = 0;
    uint64_t c[4] = {0, 0, 0, 0};
    int st = orc_align_pair(mode, qs + qoff[i], n, ts + toff[i], m, del, ext, matrix, V,
                            nullptr, nullptr, q_aln_all ? q_aln_all + i * slot : nullptr,
                            t_aln_all ? t_aln_all + i * slot : nullptr, slot, &L, c, &fv, &cv);
    if (status) status[i] = st;
    if (st) worst.store(st);
    if (f) f[i] = fv;
    if (corner) corner[i] = cv;
    if (coords) std::memcpy(coords + 4 * i, c, sizeof(c));
    if (aln_len) aln_len[i] = L;
  });
  return worst.load();
}

extern "C" int orc_heuristic_batch(const uint8_t *qs, const uint64_t *qoff, const uint8_t *ts,
                                   const uint64_t *toff, const uint32_t *pair_idx,
                                   const uint32_t *mat_idx, size_t items, double del,
                                   double ext, const double *matrices, int V, double kd,
                                   double r_squared, const double *freqs, int threads,
                                   double *f, int *passes, double *out_matrices,
                                   int *status) {
  const size_t VV = (size_t)V * V;
  std::atomic<int> worst{0};
  parallel_for(items, threads, [&](size_t i) {
    size_t p = pair_idx[i], k = mat_idx[i];
    size_t n = qoff[p + 1] - qoff[p], m = toff[p + 1] - toff[p];
    double fv = 0;
    int np = 0;
    int st = orc_heuristic(qs + qoff[p], n, ts + toff[p], m, del, ext, matrices + k * VV, V,
                           kd, r_squared, freqs + p * V, nullptr, nullptr, 0, nullptr,
                           nullptr, &fv, out_matrices ? out_matrices + i * VV : nullptr, &np);
    if (status) status[i] = st;
    if (st) worst.store(st);
    if (f) f[i] = fv;
    if (passes) passes[i] = np;
  });
  return worst.load();
}
