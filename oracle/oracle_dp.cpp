// oracle_dp.cpp — TEST INFRASTRUCTURE (see oracle.h).  Literal CPU restatement of
// aligner-core/src/enums.rs:18-46 and aligner-core/src/simple/mod.rs:42-144, :168-264.
// Compile with -ffp-contract=off so no FMA contraction changes f64 results.
#include "oracle.h"

#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

extern "C" int orc_get_direction(double top, double left, double diag, double *v) {
  // enums.rs:19  f64::max(f64::max(top, left), diagonal)  (f64::max == fmax semantics)
  double mx = std::fmax(std::fmax(top, left), diag);
  *v = mx;
  if (std::fabs(mx - top) < DBL_EPSILON) return ORC_TOP;    // enums.rs:21
  if (std::fabs(mx - left) < DBL_EPSILON) return ORC_LEFT;  // enums.rs:23
  return ORC_DIAGONAL;                                      // enums.rs:26
}

extern "C" int orc_get_direction_with_beginning(double top, double left, double diag,
                                                double *v) {
  double mx = std::fmax(std::fmax(top, left), diag);  // enums.rs:35
  *v = mx;
  if (mx == 0.0) return ORC_BEGINNING;                      // enums.rs:37
  if (std::fabs(mx - top) < DBL_EPSILON) return ORC_TOP;    // enums.rs:39
  if (std::fabs(mx - left) < DBL_EPSILON) return ORC_LEFT;  // enums.rs:41
  return ORC_DIAGONAL;
}

namespace {

// Traceback shared by both modes (simple/mod.rs:99-130 and :213-245): the output is
// seeded with the pair at (start_y, start_x) BEFORE the walk, then every move pushes.
struct Trace {
  std::vector<uint8_t> q, t;
  size_t end_x, end_y;
};

template <class GetD>
Trace walk(const uint8_t *q, const uint8_t *t, size_t start_y, size_t start_x, GetD getD) {
  Trace tr;
  size_t cx = start_x, cy = start_y;
  tr.q.push_back(q[cx - 1]);
  tr.t.push_back(t[cy - 1]);
  for (;;) {
    int d = getD(cy, cx);
    if (d == ORC_BEGINNING) break;
    if (d == ORC_TOP) {
      tr.q.push_back(ORC_BLANK);
      tr.t.push_back(t[cy - 1]);
      cy -= 1;
    } else if (d == ORC_LEFT) {
      tr.q.push_back(q[cx - 1]);
      tr.t.push_back(ORC_BLANK);
      cx -= 1;
    } else {
      tr.q.push_back(q[cx - 1]);
      tr.t.push_back(t[cy - 1]);
      cx -= 1;
      cy -= 1;
    }
  }
  tr.end_x = cx;
  tr.end_y = cy;
  // simple/mod.rs:129-130 reverse
  for (size_t i = 0, j = tr.q.size(); i + 1 < j; ++i, --j) {
    std::swap(tr.q[i], tr.q[j - 1]);
    std::swap(tr.t[i], tr.t[j - 1]);
  }
  return tr;
}

int check_symbols(const uint8_t *s, size_t len, int V) {
  for (size_t i = 0; i < len; ++i)
    if ((int)s[i] >= V) return ORC_SYMBOL_OUT_OF_RANGE;
  return ORC_OK;
}

}  // namespace

extern "C" int orc_align_pair(int mode, const uint8_t *q, size_t n, const uint8_t *t,
                              size_t m, double del, double ext, const double *matrix,
                              int V, double *H_out, uint8_t *D_out, uint8_t *q_aln,
                              uint8_t *t_aln, size_t aln_cap, size_t *aln_len,
                              uint64_t *coords, double *f, double *corner) {
  if (n == 0 || m == 0) return ORC_EMPTY_INPUT;
  if (check_symbols(q, n, V) || check_symbols(t, m, V)) return ORC_SYMBOL_OUT_OF_RANGE;

  const size_t W = n + 1;
  // simple/mod.rs:53-57 / :179-183: zeros + all Beginning, allocated per call.
  std::vector<double> Hs((m + 1) * W, 0.0);
  std::vector<uint8_t> Ds((m + 1) * W, (uint8_t)ORC_BEGINNING);
  double *H = Hs.data();
  uint8_t *D = Ds.data();

  if (mode == 0) {
    for (size_t x = 1; x <= n; ++x) {  // :59-62
      H[x] = -((double)x) * del;
      D[x] = ORC_LEFT;
    }
    for (size_t y = 1; y <= m; ++y) {  // :64-67
      H[y * W] = -((double)y) * del;
      D[y * W] = ORC_TOP;
    }
    H[n] = -((double)n + 1.0) * del;      // :69 corner quirk
    H[m * W] = -((double)m + 1.0) * del;  // :70
  }

  double penalty = del;  // :72 / :185
  for (size_t x = 1; x <= n; ++x) {      // outer: query  (:74 / :187)
    const size_t qs = q[x - 1];
    for (size_t y = 1; y <= m; ++y) {    // inner: target (:75 / :188)
      const size_t ts = t[y - 1];
      double top = H[(y - 1) * W + x] - penalty;
      double left = H[y * W + (x - 1)] - penalty;
      double diag = H[(y - 1) * W + (x - 1)] + matrix[ts * (size_t)V + qs];
      double v;
      int d = mode == 0 ? orc_get_direction(top, left, diag, &v)
                        : orc_get_direction_with_beginning(top, left, diag, &v);
      penalty = (d != ORC_BEGINNING) ? ext : del;  // :88-92 / :201-205
      H[y * W + x] = v;
      D[y * W + x] = (uint8_t)d;
    }
  }

  size_t sy, sx;
  double fval;
  if (mode == 0) {
    sy = m;
    sx = n;
    fval = 0.0;  // :139
  } else {
    // ndarray-stats 0.5.0 argmax: logical (row-major) order, strictly-greater replace.
    size_t best = 0;
    for (size_t i = 1; i < (m + 1) * W; ++i)
      if (H[i] > H[best]) best = i;
    sy = best / W;
    sx = best % W;
    if (sy == 0 || sx == 0) return ORC_NO_POSITIVE_CELL;  // :214 underflow panic
    fval = H[best];  // :247 max()
  }

  Trace tr = walk(q, t, sy, sx, [&](size_t y, size_t x) { return (int)D[y * W + x]; });

  if (aln_len) *aln_len = tr.q.size();
  if (q_aln || t_aln) {
    if (tr.q.size() > aln_cap) return ORC_BUFFER_TOO_SMALL;
    if (q_aln) std::memcpy(q_aln, tr.q.data(), tr.q.size());
    if (t_aln) std::memcpy(t_aln, tr.t.data(), tr.t.size());
  }
  if (coords) {
    if (mode == 0) {  // :137
      coords[0] = 1; coords[1] = n; coords[2] = 1; coords[3] = m;
    } else {          // :254-257
      coords[0] = tr.end_x + 1; coords[1] = sx + 1;
      coords[2] = tr.end_y + 1; coords[3] = sy + 1;
    }
  }
  if (f) *f = fval;
  if (corner) *corner = H[m * W + n];
  if (H_out) std::memcpy(H_out, H, sizeof(double) * (m + 1) * W);
  if (D_out) std::memcpy(D_out, D, (m + 1) * W);
  return ORC_OK;
}
