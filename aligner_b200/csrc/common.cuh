// common.cuh — shared device/host definitions for the alignment kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace alnr {

enum : int { DIR_TOP = 0, DIR_LEFT = 1, DIR_DIAGONAL = 2, DIR_BEGINNING = 3 };
constexpr int kBlank = 98;

// ---------------------------------------------------------------------------------------
// Direction scratch ("D") layout in HBM — shared by every fill kernel and the traceback.
// Pairs of a chunk are numbered by slot s = 0..chunk-1; 32 consecutive slots form a group
// so that the 32 lanes of a warp store one 128-byte line.  A cell (y, x), y,x >= 1, lives
// in 32-bit word (strip, x-1) of its pair, strip = (y-1) / rows_per_word, at bit
// 2 * ((y-1) % rows_per_word).  rows_per_word <= 16 (2 bits per cell, 0.25 B/cell).
//   word index = ((group * nstrips + strip) * ncols + (x-1)) * 32 + lane
// nstrips / ncols are the chunk-wide maxima.  Row 0 and column 0 are never stored: their
// directions are fixed by the mode (simple/mod.rs:59-67 global, :181-183 local).
// ---------------------------------------------------------------------------------------
//
// Two address maps share that word numbering:
//  * interleaved (above): used by the one-pair-per-thread exact kernel, whose 32 lanes store
//    word (strip, x) of 32 different pairs in the same instruction -> one 128-byte line;
//  * contiguous: word index = (slot * nstrips + strip) * ncols + (x-1), ncols a multiple of 4.
//    Used by the strip kernel, which stores 4 columns of one pair as one 16-byte vector, so a
//    traceback that moves along x reads 8 consecutive cells' words from one 32-byte sector.
struct DLayout {
  uint32_t rows_per_word;
  uint32_t nstrips;
  uint32_t ncols;
  uint32_t contiguous;
  // Bit format inside a word.  0: code of row r at bits [2r, 2r+1].  1 ("planes", strip
  // kernel): bit r = code bit 0, bit 16+r = code bit 1 — the kernel builds whole bit-planes.
  uint32_t planes;
  // Address map 3, "skewed" (warp-per-pair kernel, kernels_warp.cuh): strip = band * 32 + lane,
  // lane l works l columns behind lane 0, so the words a warp emits in one step are
  //   slot * nstrips * (ncols + 31) + ((band * (ncols + 31) + (x-1) + lane) * 32 + lane)
  // -> one 128-byte line per step.  rows_per_word = rows per lane.
  uint32_t skewed;
  // Address map 4, "blocked" (band pipeline kernel, kernels_band.cuh; single pair): a word holds
  // R = rows_per_word rows x C = 16 / R consecutive columns (blocked = C, a power of two): column
  // (x-1) % C sits at bits [2R * ((x-1) % C), ...), row r at bits [2r, 2r+1] of that field; word
  // index = strip * ncols + (x-1) / C with strip = (y-1) / R and ncols = words per strip.  A lane
  // that owns R rows packs C columns into one aligned 32-bit store.
  uint32_t blocked;
  __host__ __device__ inline uint32_t put(uint32_t code, uint32_t r) const {
    return planes ? (((code & 1u) << r) | ((code >> 1) << (16 + r))) : (code << (2 * r));
  }
  __host__ __device__ inline uint32_t get(uint32_t w, uint32_t r, uint32_t xm1 = 0) const {
    if (blocked) return (w >> (2 * rows_per_word * (xm1 & (blocked - 1)) + 2 * r)) & 3u;
    return planes ? (((w >> r) & 1u) | (((w >> (16 + r)) & 1u) << 1)) : ((w >> (2 * r)) & 3u);
  }
  __host__ __device__ inline uint64_t word(uint32_t slot, uint32_t strip, uint32_t xm1) const {
    if (blocked) return ((uint64_t)slot * nstrips + strip) * ncols + xm1 / blocked;
    if (skewed) {
      const uint32_t band = strip >> 5, lane = strip & 31;
      return (uint64_t)slot * nstrips * (ncols + 31) + (((uint64_t)band * (ncols + 31) + xm1 + lane) * 32ull + lane);
    }
    if (contiguous) return ((uint64_t)slot * nstrips + strip) * ncols + xm1;
    uint32_t group = slot >> 5, lane = slot & 31;
    return (((uint64_t)group * nstrips + strip) * ncols + xm1) * 32ull + lane;
  }
  __host__ __device__ inline uint64_t words_for(uint64_t slots) const {
    if (blocked) return slots * (uint64_t)nstrips * ncols + 4;
    if (skewed) return slots * (uint64_t)nstrips * (ncols + 31);
    if (contiguous) return slots * (uint64_t)nstrips * ncols + 4;
    return ((slots + 31) / 32) * (uint64_t)nstrips * ncols * 32ull;
  }
  static inline DLayout make(uint32_t rows_per_word, uint32_t m_max, uint32_t n_max, bool contig,
                             int planes_opt = -1) {
    DLayout d;
    d.rows_per_word = rows_per_word;
    d.nstrips = (m_max + rows_per_word - 1) / rows_per_word;
    d.ncols = contig ? (n_max + 3) / 4 * 4 : n_max;
    d.contiguous = contig ? 1u : 0u;
    d.planes = planes_opt >= 0 ? (uint32_t)planes_opt : (contig ? 1u : 0u);
    d.skewed = 0;
    d.blocked = 0;
    return d;
  }
  // band pipeline layout: R rows x 16/R columns per word (R = 1, 2, 4 or 8)
  static inline DLayout make_blocked(uint32_t R, uint32_t m_max, uint32_t n_max) {
    DLayout d;
    const uint32_t C = 16 / R;
    d.rows_per_word = R;
    d.nstrips = (m_max + R - 1) / R;
    d.ncols = ((n_max + C - 1) / C + 3) / 4 * 4;  // multiple of 4: strips start on 16-byte boundaries
    d.contiguous = 1;
    d.planes = 0;
    d.skewed = 0;
    d.blocked = C;
    return d;
  }
  // warp-per-pair layout: R rows per lane, bands of 32 * R rows
  static inline DLayout make_skewed(uint32_t R, uint32_t m_max, uint32_t n_max) {
    DLayout d;
    d.rows_per_word = R;
    d.nstrips = (m_max + 32 * R - 1) / (32 * R) * 32;
    d.ncols = n_max;
    d.contiguous = 0;
    d.planes = 0;
    d.skewed = 1;
    d.blocked = 0;
    return d;
  }
};

// Per-chunk view of the batch that every kernel receives.
struct ChunkView {
  const uint8_t *q;        // concatenated query codes (device)
  const uint8_t *t;        // concatenated target codes
  const uint64_t *q_off;   // count+1
  const uint64_t *t_off;   // count+1
  const uint32_t *order;   // slot -> pair index (nullptr = identity), whole batch
  uint32_t first;          // first slot of this chunk in `order`
  uint32_t count;          // slots in this chunk
  // Optional: result index ("pair") -> sequence index.  The heuristic search aligns the same
  // sequence pair under many matrices, so several result slots share one pair of sequences.
  const uint32_t *seqmap = nullptr;
  __device__ __forceinline__ uint32_t seq(uint32_t pair) const { return seqmap ? seqmap[pair] : pair; }
  // PWM windows (pwm/mod.rs:29-126; latent-repeat-search/engine/calc.rs:47-69): pwm_len = L > 0
  // switches the view: item i is the window t[win_start(i), +win_len(i)) of ONE resident sequence
  // (the DP rows) against the L columns of a 4 x L profile; there is no query sequence and the
  // column "symbol" of column x is x - 1.  Windows are explicit (win_starts / win_lens) or strided.
  uint32_t pwm_len = 0;
  const uint64_t *win_starts = nullptr;
  const uint32_t *win_lens = nullptr;
  uint64_t win_first = 0, win_step = 0, seq_len = 0;
  uint32_t win_size = 0;
  __device__ __forceinline__ void pair_dims(uint32_t sp, const uint8_t *&qp, const uint8_t *&tp,
                                            uint32_t &n, uint32_t &m) const {
    if (pwm_len) {
      uint64_t s0;
      if (win_starts) { s0 = win_starts[sp]; m = win_lens[sp]; }
      else {
        s0 = win_first + (uint64_t)sp * win_step;
        m = s0 >= seq_len ? 0u : (uint32_t)(seq_len - s0 < win_size ? seq_len - s0 : win_size);
      }
      qp = nullptr; tp = t + s0; n = pwm_len;
    } else {
      qp = q + q_off[sp]; tp = t + t_off[sp];
      n = (uint32_t)(q_off[sp + 1] - q_off[sp]);
      m = (uint32_t)(t_off[sp + 1] - t_off[sp]);
    }
  }
};

// Per-pair results produced by the fill kernels and consumed by the traceback.
struct FillOut {
  double *f;        // local: max(H); global: 0.0            [pair]
  double *corner;   // H[m][n]                               [pair]
  uint32_t *start;  // traceback start (y, x)                [pair*2]
  int32_t *status;  // alnr_status per pair                  [pair]
};

struct TraceOut {
  uint64_t *coords;   // [pair*4]
  uint64_t *aln_off;  // [pair]
  uint32_t *aln_len;  // [pair]
  uint8_t *q_aln;     // slots, see aligner_b200.h
  uint8_t *t_aln;
  double *freq;       // optional V*V count matrix per pair (alignment.rs:13-23), or nullptr
};

}  // namespace alnr
