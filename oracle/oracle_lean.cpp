// oracle_lean.cpp — TEST INFRASTRUCTURE (see oracle.h).  Same recurrence as
// oracle_dp.cpp (simple/mod.rs:42-144, :168-264) with a rolling score column and 2-bit
// directions, so that 10^5 x 10^5 pairs fit in host memory.  Results are identical by
// construction; tests/ checks lean == faithful on small inputs.
#include "oracle.h"

#include <cstdio>
#include <cstring>
#include <vector>

namespace {

struct Sha256 {
  uint32_t h[8];
  uint8_t buf[64];
  size_t fill = 0;
  uint64_t total = 0;
  Sha256() {
    static const uint32_t init[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a,
                                     0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    std::memcpy(h, init, sizeof(h));
  }
  static uint32_t rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
  void block(const uint8_t *p) {
    static const uint32_t K[64] = {
        0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4,
        0xab1c5ed5, 0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe,
        0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f,
        0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7,
        0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc,
        0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b,
        0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116,
        0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
        0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7,
        0xc67178f2};
    uint32_t w[64];
    for (int i = 0; i < 16; ++i)
      w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 |
             (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
    for (int i = 16; i < 64; ++i) {
      uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3);
      uint32_t s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
      w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    for (int i = 0; i < 64; ++i) {
      uint32_t S1 = rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25);
      uint32_t ch = (e & f) ^ (~e & g);
      uint32_t t1 = hh + S1 + ch + K[i] + w[i];
      uint32_t S0 = rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22);
      uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
      uint32_t t2 = S0 + mj;
      hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
  }
  void update(const uint8_t *p, size_t len) {
    total += len;
    while (len) {
      size_t k = 64 - fill < len ? 64 - fill : len;
      std::memcpy(buf + fill, p, k);
      fill += k; p += k; len -= k;
      if (fill == 64) { block(buf); fill = 0; }
    }
  }
  void hex16(char *out) {
    uint64_t bits = total * 8;
    uint8_t pad = 0x80;
    update(&pad, 1);
    uint8_t z = 0;
    while (fill != 56) update(&z, 1);
    uint8_t lenb[8];
    for (int i = 0; i < 8; ++i) lenb[i] = (uint8_t)(bits >> (56 - 8 * i));
    update(lenb, 8);
    std::snprintf(out, 17, "%08x%08x", h[0], h[1]);
  }
};

struct Bits2 {  // 2-bit cell store, row-major [(m+1)][(n+1)]
  std::vector<uint8_t> v;
  explicit Bits2(size_t cells) : v((cells + 3) / 4, 0) {}
  void set(size_t i, int d) { v[i >> 2] |= (uint8_t)(d << ((i & 3) * 2)); }
  int get(size_t i) const { return (v[i >> 2] >> ((i & 3) * 2)) & 3; }
};

}  // namespace

extern "C" int orc_align_pair_lean(int mode, const uint8_t *q, size_t n, const uint8_t *t,
                                   size_t m, double del, double ext, const double *matrix,
                                   int V, uint8_t *q_aln, uint8_t *t_aln, size_t aln_cap,
                                   size_t *aln_len, uint64_t *coords, double *f,
                                   double *corner, char *d_sha_hex) {
  if (n == 0 || m == 0) return ORC_EMPTY_INPUT;
  for (size_t i = 0; i < n; ++i) if ((int)q[i] >= V) return ORC_SYMBOL_OUT_OF_RANGE;
  for (size_t i = 0; i < m; ++i) if ((int)t[i] >= V) return ORC_SYMBOL_OUT_OF_RANGE;
  const size_t W = n + 1;
  Bits2 D((m + 1) * W);
  std::vector<double> prev(m + 1), cur(m + 1);  // columns x-1 and x

  // column 0
  for (size_t y = 0; y <= m; ++y) {
    if (mode == 0) {
      prev[y] = -((double)y) * del;
      D.set(y * W, y == 0 ? ORC_BEGINNING : ORC_TOP);
    } else {
      prev[y] = 0.0;
      D.set(y * W, ORC_BEGINNING);
    }
  }
  if (mode == 0) prev[m] = -((double)m + 1.0) * del;  // simple/mod.rs:70
  prev[0] = 0.0;

  double penalty = del;
  double best = 0.0;  // H[0][0]; row-major-first arg-max (smallest y, then smallest x)
  size_t by = 0, bx = 0;
  for (size_t x = 1; x <= n; ++x) {
    if (mode == 0) {
      cur[0] = (x == n) ? -((double)n + 1.0) * del : -((double)x) * del;  // :59-62, :69
      D.set(x, ORC_LEFT);
    } else {
      cur[0] = 0.0;
      D.set(x, ORC_BEGINNING);
    }
    const size_t qs = q[x - 1];
    for (size_t y = 1; y <= m; ++y) {
      double top = cur[y - 1] - penalty;
      double left = prev[y] - penalty;
      double diag = prev[y - 1] + matrix[(size_t)t[y - 1] * (size_t)V + qs];
      double v;
      int d = mode == 0 ? orc_get_direction(top, left, diag, &v)
                        : orc_get_direction_with_beginning(top, left, diag, &v);
      penalty = (d != ORC_BEGINNING) ? ext : del;
      cur[y] = v;
      D.set(y * W + x, d);
      if (mode == 1 && (v > best || (v == best && (y < by || (y == by && x < bx))))) {
        best = v; by = y; bx = x;
      }
    }
    prev.swap(cur);
  }
  if (corner) *corner = prev[m];

  size_t sy = m, sx = n;
  double fval = 0.0;
  if (mode == 1) {
    if (by == 0 || bx == 0) return ORC_NO_POSITIVE_CELL;
    sy = by; sx = bx; fval = best;
  }

  std::vector<uint8_t> qa, ta;
  size_t cx = sx, cy = sy;
  qa.push_back(q[cx - 1]);
  ta.push_back(t[cy - 1]);
  for (;;) {
    int d = D.get(cy * W + cx);
    if (d == ORC_BEGINNING) break;
    if (d == ORC_TOP) { qa.push_back(ORC_BLANK); ta.push_back(t[cy - 1]); cy--; }
    else if (d == ORC_LEFT) { qa.push_back(q[cx - 1]); ta.push_back(ORC_BLANK); cx--; }
    else { qa.push_back(q[cx - 1]); ta.push_back(t[cy - 1]); cx--; cy--; }
  }
  const size_t L = qa.size();
  if (aln_len) *aln_len = L;
  if (q_aln || t_aln) {
    if (L > aln_cap) return ORC_BUFFER_TOO_SMALL;
    for (size_t i = 0; i < L; ++i) {
      if (q_aln) q_aln[i] = qa[L - 1 - i];
      if (t_aln) t_aln[i] = ta[L - 1 - i];
    }
  }
  if (coords) {
    if (mode == 0) { coords[0] = 1; coords[1] = n; coords[2] = 1; coords[3] = m; }
    else { coords[0] = cx + 1; coords[1] = sx + 1; coords[2] = cy + 1; coords[3] = sy + 1; }
  }
  if (f) *f = fval;
  if (d_sha_hex) {
    Sha256 sha;
    std::vector<uint8_t> row(W);
    for (size_t y = 0; y <= m; ++y) {
      for (size_t x = 0; x <= n; ++x) row[x] = (uint8_t)D.get(y * W + x);
      sha.update(row.data(), W);
    }
    sha.hex16(d_sha_hex);
  }
  return ORC_OK;
}
