// oracle_pwm.cpp — TEST INFRASTRUCTURE (see oracle.h).  Literal CPU restatement of
// aligner-core/src/pwm/mod.rs:29-126 (PWMAligner::perform_alignment) and of
// PWMAlignment::get_frequency_matrix / get_alignment (aligner-core/src/alignment.rs:54-79).
//
// PARITY UNPINNED for this row: the reference holds no test, fixture or golden vector for the
// PWM aligner and cannot be built here (Rust).  The direction / tie / penalty
// functions it calls are the ones oracle_dp.cpp pins against the reference's golden vectors.
// Compile with -ffp-contract=off.
#include "oracle.h"

#include <cstring>
#include <thread>
#include <vector>

extern "C" int orc_pwm_align(const uint8_t *query, size_t m, const double *pwm, size_t rows,
                             size_t L, double del, double ext, double *H_out, uint8_t *D_out,
                             uint32_t *numbered, uint8_t *q_aln, size_t aln_cap, size_t *aln_len,
                             uint64_t *coords, double *f) {
  if (rows != 4) return ORC_MATRIX_SHAPE_ERROR;  // pwm/mod.rs:40-42
  for (size_t i = 0; i < m; ++i)
    if (query[i] >= 4) return ORC_SYMBOL_OUT_OF_RANGE;  // matrix[[seq_2_pos, ..]] bounds panic
  // numbered_sequence = 1..=L (:44); dim = (query.len() + 1, L + 1) (:46)
  const size_t W = L + 1;
  std::vector<double> H((m + 1) * W, 0.0);                    // :48
  std::vector<uint8_t> D((m + 1) * W, (uint8_t)ORC_BEGINNING);  // :49-50
  double penalty = del;                                       // :52
  for (size_t x = 1; x <= L; ++x) {                           // :54 elem_1
    for (size_t y = 0; y < m; ++y) {                          // :55
      const size_t yr = y + 1;
      const size_t sym = query[y];
      double v;
      const int d = orc_get_direction_with_beginning(H[(yr - 1) * W + x] - penalty,       // :60
                                                     H[yr * W + (x - 1)] - penalty,       // :61
                                                     H[(yr - 1) * W + (x - 1)] + pwm[sym * L + (x - 1)],  // :62
                                                     &v);
      penalty = (d != ORC_BEGINNING) ? ext : del;             // :65-69
      H[yr * W + x] = v;                                      // :71
      D[yr * W + x] = (uint8_t)d;                             // :72
    }
  }
  // :76 argmax (ndarray-stats 0.5.0: row-major, strictly-greater replace => first maximum)
  size_t best = 0;
  for (size_t i = 1; i < (m + 1) * W; ++i)
    if (H[i] > H[best]) best = i;
  const size_t my = best / W, mx = best % W;
  size_t cx = mx, cy = my;
  std::vector<uint32_t> num;
  std::vector<uint8_t> qa;
  for (;;) {                                                  // :81-104, no seed push
    const int d = D[cy * W + cx];
    if (d == ORC_BEGINNING) break;
    if (d == ORC_TOP) { num.push_back(0); qa.push_back(query[cy - 1]); cy -= 1; }
    else if (d == ORC_LEFT) { num.push_back((uint32_t)cx); qa.push_back(ORC_BLANK); cx -= 1; }
    else { num.push_back((uint32_t)cx); qa.push_back(query[cy - 1]); cx -= 1; cy -= 1; }
  }
  const size_t len = num.size();
  if (aln_len) *aln_len = len;
  if (numbered || q_aln) {
    if (len > aln_cap) return ORC_BUFFER_TOO_SMALL;
    for (size_t i = 0; i < len; ++i) {                        // :106-107 reverse
      if (numbered) numbered[i] = num[len - 1 - i];
      if (q_aln) q_aln[i] = qa[len - 1 - i];
    }
  }
  if (f) *f = H[best];                                        // :109
  if (coords) {                                               // :118-121
    coords[0] = cx + 1; coords[1] = mx + 1; coords[2] = cy + 1; coords[3] = my + 1;
  }
  if (H_out) std::memcpy(H_out, H.data(), sizeof(double) * H.size());
  if (D_out) std::memcpy(D_out, D.data(), D.size());
  return ORC_OK;
}

// alignment.rs:55-65: counts[[sym, pos - 1]] over columns with pos != 0 and sym != blank
extern "C" void orc_pwm_frequency_matrix(const uint32_t *numbered, const uint8_t *q_aln, size_t len,
                                         size_t L, double *out /* 4*L */) {
  for (size_t i = 0; i < 4 * L; ++i) out[i] = 0.0;
  for (size_t i = 0; i < len; ++i)
    if (numbered[i] != 0 && q_aln[i] != ORC_BLANK) out[(size_t)q_aln[i] * L + numbered[i] - 1] += 1.0;
}

// Windows of one sequence against one PWM (latent-repeat-search/engine/calc.rs:47-69, :107-137):
// window i = seq[starts[i], starts[i] + lens[i]).  Alignments go to slots of `slot` entries.
extern "C" void orc_pwm_windows(const uint8_t *seq, const uint64_t *starts, const uint32_t *lens,
                                size_t count, const double *pwm, size_t L, double del, double ext,
                                int threads, double *f, uint64_t *coords, uint64_t *aln_len,
                                uint32_t *numbered, uint8_t *q_aln, size_t slot, int32_t *status) {
  auto work = [&](size_t lo, size_t hi) {
    for (size_t i = lo; i < hi; ++i) {
      size_t len = 0;
      status[i] = orc_pwm_align(seq + starts[i], lens[i], pwm, 4, L, del, ext, nullptr, nullptr,
                                numbered ? numbered + i * slot : nullptr, q_aln ? q_aln + i * slot : nullptr,
                                slot, &len, coords ? coords + 4 * i : nullptr, f ? f + i : nullptr);
      if (aln_len) aln_len[i] = len;
    }
  };
  if (threads <= 1) { work(0, count); return; }
  std::vector<std::thread> th;
  for (int k = 0; k < threads; ++k) th.emplace_back(work, count * k / threads, count * (k + 1) / threads);
  for (auto &t : th) t.join();
}
