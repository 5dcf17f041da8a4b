// expand.cu — host side of the compact result format (alnr_batch_out.moves): the reference's two
// aligned code strings, coords and frequency matrices rebuilt from the 2-bit move strings.
//
// The device stops the traceback at the walk itself (start cell + one Direction per step,
// simple/mod.rs:107-127, :220-242).  What the reference's Alignment holds is a pure function of that
// walk and the two input sequences:
//   column aln_len-1 (the seed, :102-105 / :213-216) = (q[sx-1], t[sy-1]);
//   step k lands in column aln_len-2-k: Top -> (Blank, t[cy-1]), cy--; Left -> (q[cx-1], Blank),
//   cx--; Diagonal -> (q[cx-1], t[cy-1]), both-- (:110-124, :227-241);
//   coords = ((cx+1, sx+1), (cy+1, sy+1)) in local mode (:254-257), ((1, n), (1, m)) in global (:137).
// No CUDA in this file: it runs on the caller's host threads.
#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/aligner_b200.h"

namespace {

struct SeqView {
  const uint8_t *bytes;   // byte codes, or
  const uint32_t *words;  // 2-bit packed words of THIS sequence
  inline uint8_t at(uint64_t k) const {
    return bytes ? bytes[k] : (uint8_t)((words[k >> 4] >> (2 * (k & 15))) & 3u);
  }
};

}  // namespace

extern "C" int alnr_expand_alignments(int mode, int volume, const alnr_batch *b, const alnr_batch_out *c,
                                      uint64_t first, uint64_t count, alnr_batch_out *ex, double *freq,
                                      int threads) {
  if (!b || !c || !ex || !c->moves || !c->aln_len || !c->end_yx) return ALNR_INVALID_ARGUMENT;
  if (!b->q_off || !b->t_off || first + count > b->count) return ALNR_INVALID_ARGUMENT;
  if (!c->moves_off && first != 0) return ALNR_INVALID_ARGUMENT;
  if ((ex->q_aln != nullptr) != (ex->t_aln != nullptr)) return ALNR_INVALID_ARGUMENT;
  if (freq && (volume < 1 || volume > 32)) return ALNR_MATRIX_SHAPE_ERROR;
  const bool packed = b->q2 != nullptr && b->t2 != nullptr;
  if (!packed && (!b->q || !b->t)) return ALNR_INVALID_ARGUMENT;
  if (count == 0) return ALNR_OK;
  try {
  unsigned nt = threads > 0 ? (unsigned)threads : std::max(1u, std::thread::hardware_concurrency());
  nt = (unsigned)std::min<uint64_t>(nt, (count + 4095) / 4096);
  if (nt < 1) nt = 1;

  // ---- pass 1 (per part): bytes of the strings, move words, packed-input words -----------------
  // Offsets of pair i within the range: alignment bytes before it, and — when the caller kept no
  // moves_off — move words before it; for 2-bit inputs the word offset of its sequences.
  std::vector<uint64_t> part_bytes(nt + 1, 0), part_mw(nt + 1, 0), part_qw(nt + 1, 0), part_tw(nt + 1, 0);
  auto bounds = [&](unsigned k, uint64_t &lo, uint64_t &hi) { lo = first + count * k / nt; hi = first + count * (k + 1) / nt; };
  auto seq_words = [&](const uint64_t *off, uint64_t i) { return (off[i + 1] - off[i] + 15) >> 4; };
  auto pass1 = [&](unsigned k) {
    uint64_t lo, hi, by = 0, mw = 0, qw = 0, tw = 0;
    bounds(k, lo, hi);
    for (uint64_t i = lo; i < hi; ++i) {
      by += c->aln_len[i];
      mw += ((uint64_t)c->aln_len[i] + 14) >> 4;
      if (packed) { qw += seq_words(b->q_off, i); tw += seq_words(b->t_off, i); }
    }
    part_bytes[k + 1] = by; part_mw[k + 1] = mw; part_qw[k + 1] = qw; part_tw[k + 1] = tw;
  };
  // 2-bit inputs: words of the sequences before `first` (dense packing, see aligner_b200.h)
  uint64_t qw0 = 0, tw0 = 0;
  if (packed)
    for (uint64_t i = 0; i < first; ++i) { qw0 += seq_words(b->q_off, i); tw0 += seq_words(b->t_off, i); }
  {
    std::vector<std::thread> th;
    for (unsigned k = 1; k < nt; ++k) th.emplace_back(pass1, k);
    pass1(0);
    for (auto &t : th) t.join();
  }
  part_qw[0] = qw0; part_tw[0] = tw0;
  for (unsigned k = 0; k < nt; ++k) {
    part_bytes[k + 1] += part_bytes[k]; part_mw[k + 1] += part_mw[k];
    part_qw[k + 1] += part_qw[k]; part_tw[k + 1] += part_tw[k];
  }

  // ---- pass 2: expand -------------------------------------------------------------------------
  const bool glob = mode == ALNR_GLOBAL;
  const int V = volume;
  auto pass2 = [&](unsigned k) {
    uint64_t lo, hi;
    bounds(k, lo, hi);
    uint64_t ob = part_bytes[k], mw = part_mw[k], qw = part_qw[k], tw = part_tw[k];
    for (uint64_t i = lo; i < hi; ++i) {
      const uint64_t r = i - first;
      const uint32_t L = c->aln_len[i];
      const uint64_t n = b->q_off[i + 1] - b->q_off[i], m = b->t_off[i + 1] - b->t_off[i];
      SeqView q{packed ? nullptr : b->q + b->q_off[i], packed ? b->q2 + qw : nullptr};
      SeqView t{packed ? nullptr : b->t + b->t_off[i], packed ? b->t2 + tw : nullptr};
      if (packed) { qw += (n + 15) >> 4; tw += (m + 15) >> 4; }
      const uint32_t *mv = c->moves + (c->moves_off ? c->moves_off[i] : mw);
      mw += ((uint64_t)L + 14) >> 4;
      if (ex->aln_len) ex->aln_len[r] = L;
      if (ex->aln_off) ex->aln_off[r] = ob;
      double *fm = freq ? freq + r * (uint64_t)V * V : nullptr;
      if (fm) std::fill(fm, fm + (size_t)V * V, 0.0);
      if (L == 0) {  // a pair with a status owns no alignment
        if (ex->coords) std::memset(ex->coords + 4 * r, 0, 32);
        continue;
      }
      const uint32_t sy = c->end_yx[2 * i], sx = c->end_yx[2 * i + 1];
      uint64_t cy = sy, cx = sx;
      uint8_t *qa = ex->q_aln ? ex->q_aln + ob : nullptr, *ta = ex->t_aln ? ex->t_aln + ob : nullptr;
      auto emit = [&](uint64_t pos, uint8_t qc, uint8_t tc) {
        if (qa) { qa[pos] = qc; ta[pos] = tc; }
        if (fm && qc != ALNR_BLANK && tc != ALNR_BLANK) fm[(size_t)tc * V + qc] += 1.0;  // alignment.rs:16-20
      };
      emit(L - 1, q.at(cx - 1), t.at(cy - 1));
      uint32_t word = 0;
      for (uint32_t s = 0; s + 1 < L; ++s) {
        if ((s & 15) == 0) word = mv[s >> 4];
        const uint32_t d = word & 3u;
        word >>= 2;
        const uint64_t pos = L - 2 - s;
        if (d == ALNR_DIR_TOP) { emit(pos, ALNR_BLANK, t.at(cy - 1)); --cy; }
        else if (d == ALNR_DIR_LEFT) { emit(pos, q.at(cx - 1), ALNR_BLANK); --cx; }
        else { emit(pos, q.at(cx - 1), t.at(cy - 1)); --cx; --cy; }
      }
      if (ex->coords) {
        uint64_t *co = ex->coords + 4 * r;
        if (glob) { co[0] = 1; co[1] = n; co[2] = 1; co[3] = m; }
        else { co[0] = cx + 1; co[1] = (uint64_t)sx + 1; co[2] = cy + 1; co[3] = (uint64_t)sy + 1; }
      }
      ob += L;
    }
  };
  {
    std::vector<std::thread> th;
    for (unsigned k = 1; k < nt; ++k) th.emplace_back(pass2, k);
    pass2(0);
    for (auto &t : th) t.join();
  }
  } catch (...) {  // std::thread / std::vector: nothing may cross the C boundary
    return ALNR_OUT_OF_MEMORY;
  }
  return ALNR_OK;
}
