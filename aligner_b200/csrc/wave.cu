// wave.cu — host side of the intra-pair wavefront path for one long pair: the tile pipeline
// (kernels_tile.cuh) and the checkpointed traceback.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "kernels_diff.cuh"
#include "kernels_tile.cuh"
#include "kernels_trace.cuh"
#include "strip_launch.h"
#include "wave_launch.h"

namespace alnr {

// One pair at a time on the tile pipeline pays off from ~3 x 10^4 cells for single-pair calls
// (scripts/pair_sizes.py, global DNA pairs: 200 x 200 0.22 against 0.29 ms on the warp kernel,
// 600 x 600 0.29 against 0.74 ms, 1000 x 1000 0.35 against 1.22 ms); in a batch the warp-per-pair
// kernel runs many pairs side by side, so only very long pairs qualify.  ALNR_WAVE_MIN_CELLS
// overrides the single-call threshold (A/B runs).
static const uint64_t kWaveMinCellsSingle =
    std::getenv("ALNR_WAVE_MIN_CELLS") ? std::strtoull(std::getenv("ALNR_WAVE_MIN_CELLS"), nullptr, 10) : 1ull << 15;
static const uint64_t kWaveMinCellsBatch = 1ull << 24;

bool wave_eligible(const alnr_batch_impl &B, uint32_t m, uint32_t n) {
  if (B.numeric != ALNR_NUM_I32 || B.want_matrices) return false;
  if ((uint64_t)m * n < (B.count <= 4 ? kWaveMinCellsSingle : kWaveMinCellsBatch)) return false;
  if (m >= 0xFFFFFu || n >= 0xFFFFFu) return false;       // arg-max key packs 20-bit coordinates
  // stateful penalty (local, del != ext): the tile pipeline speculates row 1's penalties and repeats
  // the pass until they reproduce themselves (tile_run_pair); worth it from ~2000 x 2000 cells on
  if (B.mode == ALNR_LOCAL && B.del != B.ext && ((uint64_t)m * n < (1ull << 22) || std::getenv("ALNR_NO_WAVE_STATEFUL")))
    return false;
  double mag = 0;
  for (double v : B.matrix) mag = std::max(mag, v);
  if (B.mode == ALNR_LOCAL && mag * std::min(m, n) >= 16777216.0) return false;  // 24-bit value field
  return true;
}

// Bands per group of the checkpointed traceback for a pair of m x n cells, or 0 when the whole
// direction matrix (m * n / 4 bytes) fits `budget` bytes.  ALNR_SCRATCH_BUDGET (bytes) overrides the
// budget (tests).
uint32_t wave_ckpt_group(uint32_t m, uint32_t n, uint64_t budget) {
  if (const char *e = std::getenv("ALNR_SCRATCH_BUDGET")) {  // read per call: tests switch it on and off
    const uint64_t forced = std::strtoull(e, nullptr, 10);
    if (forced) budget = forced;
  }
  const DLayout full = DLayout::make_blocked(4, m, n);
  if (full.words_for(1) * 4 <= budget) return 0;
  const uint64_t per_band = 32ull * full.ncols * 4;  // 32 strips of one 128-row band
  const uint32_t nbands = (m + 127) / 128;
  return (uint32_t)std::min<uint64_t>(std::max<uint64_t>(budget / per_band, 1), nbands);
}

static int tile_args(alnr_batch_impl &B, const Chunk &c, BandArgs &a, size_t &smem, unsigned &resident);

// Checkpointed traceback (BASELINE north_star (2): "checkpointed Hirschberg-style for very long
// pairs"; the loop it reproduces: simple/mod.rs:99-130, :213-245).  The first pass ran score-only
// and kept the bottom row of every 128-row band (m * n / 32 bytes instead of m * n / 4).  The path
// is then walked group by group from the bottom: the group's bands are filled again WITH directions
// into a scratch of `group` bands — every band starts from the kept row above it, so the bands of a
// group run side by side — and the walk crosses the group and parks at its top row.  Everything is
// enqueued without host synchronisation: a re-fill whose group the path does not cross (local
// mode: the alignment starts and ends anywhere) returns at once.
int wave_ckpt_trace(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const TraceArgs &ta_in, cudaStream_t s) {
  BandArgs a;
  size_t smem = 0;
  unsigned resident = 1;
  int st = tile_args(B, c, a, smem, resident);
  if (st) return st;
  const uint32_t G = c.ckpt_group, nbands = a.nbands;
  if (!B.d_ckpt.p) ALNR_CUDA(B.d_ckpt.alloc(8));
  CkptState *ck = reinterpret_cast<CkptState *>(B.d_ckpt.p);
  TraceArgs ta = ta_in;
  ta.dl = dl;
  ckpt_init_kernel<<<1, 1, 0, s>>>(ta, ck);
  ALNR_LAUNCHED();
  const bool local = B.mode == ALNR_LOCAL;
  a.best = nullptr; a.corner_out = nullptr; a.ckpt = B.d_ckpt.p;
  a.D = B.Dp();
  for (uint32_t b1 = nbands; b1 > 0;) {
    const uint32_t b0 = b1 > G ? b1 - G : 0;
    a.band0 = b0; a.band_end = b1; a.strip0 = b0 * 32;
    const unsigned grid = std::min<unsigned>(b1 - b0, resident);
    if (local) fill_tile_kernel<true, true><<<grid, 32, smem, s>>>(a);
    else fill_tile_kernel<false, true><<<grid, 32, smem, s>>>(a);
    ALNR_LAUNCHED();
    traceback_warp_kernel<<<1, 32, 0, s>>>(ta, ck, b0 * 128u, b0 * 32u);
    ALNR_LAUNCHED();
    b1 = b0;
  }
  // the border run of a global path (column 0 / row 0 need no directions) and the result record
  traceback_warp_kernel<<<1, 32, 0, s>>>(ta, ck, 0u, 0u);
  ALNR_LAUNCHED();
  return ALNR_OK;
}

// BandArgs, shared-memory size and resident grid of the tile kernel for the pair of chunk c
static int tile_args(alnr_batch_impl &B, const Chunk &c, BandArgs &a, size_t &smem, unsigned &resident) {
  a.q = B.d_q.p + c.q0; a.t = B.d_t.p + c.t0;
  a.n = c.n_max; a.m = c.m_max; a.V = B.V; a.matrix = B.d_mati.p;
  a.del = (int)B.del; a.ext = (int)B.ext;
  a.nbands = (a.m + 127) / 128;
  a.prog = nullptr;
  a.rows = B.d_band.p + 4;
  a.dcols = DLayout::make_blocked(4, a.m, a.n).ncols;
  smem = (size_t)B.V * 32 * sizeof(int4) + 32 * sizeof(int);
  int occ0 = 0, occ1 = 0, sms = 0, dev = 0;
  ALNR_CUDA(cudaGetDevice(&dev));
  ALNR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ0, fill_tile_kernel<false>, 32, smem));
  ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, fill_tile_kernel<true>, 32, smem));
  resident = (unsigned)std::max(1, std::min(occ0, occ1)) * (unsigned)std::max(1, sms);
  return ALNR_OK;
}

// Stateful local pairs: what a pass of the tile pipeline needs reset, unless an earlier pass converged.
static __global__ void stateful_reset_kernel(BandArgs a, size_t row_words, uint32_t map_words) {
  if (a.prog[0]) return;
  const size_t stride = (size_t)gridDim.x * blockDim.x, tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t *zeromap = stateful_zeromap(a);
  for (size_t i = tid; i < row_words; i += stride) a.rows[i] = kTileSentinel;
  for (size_t i = tid; i < map_words; i += stride) zeromap[i] = 0;
  if (tid == 0) { *a.best = 0; *a.corner_out = 0; }
}

// After a pass: cell (1, x) must pay del iff H[m][x-1] == 0 (simple/mod.rs:201-205), cell (1,1) pays
// del.  In map bits (bit c = column c + 1): wanted = (zeromap << 1) | 1.  Equal to the map the pass
// ran with: converged.  Otherwise the pass is repeated with the wanted map.  One block.
static __global__ void stateful_check_kernel(BandArgs a, uint32_t map_words) {
  if (a.prog[0]) return;
  uint32_t *delmap = stateful_delmap(a);
  const uint32_t *zeromap = stateful_zeromap(a);
  __shared__ int differs;
  if (threadIdx.x == 0) differs = 0;
  __syncthreads();
  const uint32_t nwords = (a.n + 31) / 32;
  for (uint32_t w = threadIdx.x; w < map_words; w += blockDim.x) {
    uint32_t want = 0;
    if (w < nwords) {
      want = (zeromap[w] << 1) | (w ? zeromap[w - 1] >> 31 : 1u);
      if (w + 1 == nwords && (a.n & 31u)) want &= (1u << (a.n & 31u)) - 1u;  // columns 1 .. n only
    }
    if (want != delmap[w]) { differs = 1; delmap[w] = want; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    a.prog[1] += 1;
    if (!differs) a.prog[0] = 1;
  }
}

// Tile pipeline (kernels_tile.cuh): 4 x 4 cells per lane and step, sentinel hand-off of the band rows.
static int tile_run_pair(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const FillOut &fo) {
  BandArgs a;
  a.q = B.d_q.p + c.q0; a.t = B.d_t.p + c.t0;
  a.n = c.n_max; a.m = c.m_max; a.V = B.V; a.matrix = B.d_mati.p;
  a.del = (int)B.del; a.ext = (int)B.ext;
  a.nbands = (a.m + 127) / 128;
  const bool local = B.mode == ALNR_LOCAL;
  const size_t npad = ((size_t)a.n + 3) / 4 * 4;
  // scratch: [best (8 B)][corner][pad][rows nbands * npad][32 ints of slack for the first prefetch]
  const size_t row_words = (size_t)a.nbands * npad + 32;
#ifdef ALNR_TILE_DEBUG
  const size_t dbg_words = (size_t)a.nbands * 16 + 4;
#else
  const size_t dbg_words = 0;
#endif
  // stateful local pair: [state 4 words][delmap][zeromap], each map (n + 31) / 32 words + 4 of slack
  const bool stateful = local && B.del != B.ext;
  const uint32_t map_words = stateful ? stateful_map_words(a.n) : 0;
  const size_t words = 4 + row_words + dbg_words + (stateful ? 4 + 2 * (size_t)map_words : 0);
  if (B.d_band.n < words) {
    // a longer pair than the previous one of this batch: its fill may still be running on the old rows
    if (B.d_band.p) ALNR_CUDA(cudaStreamSynchronize(B.stream));
    ALNR_CUDA(B.d_band.alloc(words));
  }
  int *base = B.d_band.p;
  a.best = reinterpret_cast<unsigned long long *>(base);
  a.corner_out = base + 2;
  a.prog = nullptr;
  a.rows = base + 4;
#ifdef ALNR_TILE_DEBUG
  a.prog = reinterpret_cast<uint32_t *>(base + 4 + (row_words + 3) / 4 * 4);
#endif
  a.D = c.ckpt_group ? nullptr : B.Dp();  // checkpointed: score-only first pass, directions group by group later
  a.dcols = dl.ncols;
  const size_t smem = (size_t)B.V * 32 * sizeof(int4) + 32 * sizeof(int);
  int occ0 = 0, occ1 = 0, sms = 0, dev = 0;
  ALNR_CUDA(cudaGetDevice(&dev));
  ALNR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ0, fill_tile_kernel<false>, 32, smem));
  ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, fill_tile_kernel<true>, 32, smem));
  const unsigned resident = (unsigned)std::max(1, std::min(occ0, occ1)) * (unsigned)std::max(1, sms);
  ALNR_CUDA(cudaMemsetAsync(a.best, 0, 16, B.stream));
  if (a.nbands > 1)  // every word of the boundary rows starts as "not produced yet" (0x80808080)
    ALNR_CUDA(cudaMemsetAsync(a.rows, 0x80, (size_t)(a.nbands - 1) * npad * sizeof(int), B.stream));
  const unsigned grid = std::min<unsigned>(a.nbands, resident);  // every block must be resident
  if (stateful) {
    // Speculate-and-repeat (kernels_band.cuh): kStatefulPasses passes are enqueued; a pass whose
    // penalty map reproduced itself sets state[0] and the later ones return at once, so nothing
    // synchronises with the host.  A map still moving after the last pass ends as
    // ALNR_CALCULATION_ERROR for the pair (band_finish_kernel) — every pass pins at least the first
    // column that differed, and zeros in row m are rare away from the matrix's left edge.
    constexpr int kStatefulPasses = 8;
    uint32_t *sbase = reinterpret_cast<uint32_t *>(base + 4 + row_words + dbg_words);
    uint32_t *delmap = sbase + 4;
    a.prog = sbase;  // the state block (kernels_band.cuh)
    ALNR_CUDA(cudaMemsetAsync(sbase, 0, (4 + 2 * (size_t)map_words) * sizeof(uint32_t), B.stream));
    const uint32_t one = 1;  // cell (1,1) pays del (simple/mod.rs:185)
    ALNR_CUDA(cudaMemcpyAsync(delmap, &one, sizeof(one), cudaMemcpyHostToDevice, B.stream));
    const size_t reset_words = (size_t)(a.nbands - 1) * npad;
    for (int pass = 0; pass < kStatefulPasses; ++pass) {
      if (pass) {  // the first pass starts from the memsets above
        stateful_reset_kernel<<<148 * 4, 256, 0, B.stream>>>(a, reset_words, map_words);
        ALNR_LAUNCHED();
      }
      fill_tile_kernel<true, false, true><<<grid, 32, smem, B.stream>>>(a);
      ALNR_LAUNCHED();
      stateful_check_kernel<<<1, 256, 0, B.stream>>>(a, map_words);
      ALNR_LAUNCHED();
    }
    if (std::getenv("ALNR_WAVE_STATS")) {  // tests: passes run, through alnr_batch_rerun_count (synchronises)
      uint32_t st[2] = {0, 0};
      ALNR_CUDA(cudaMemcpyAsync(st, sbase, sizeof(st), cudaMemcpyDeviceToHost, B.stream));
      ALNR_CUDA(cudaStreamSynchronize(B.stream));
      B.rerun_total += st[1];
    }
  } else if (local) fill_tile_kernel<true><<<grid, 32, smem, B.stream>>>(a);
  else fill_tile_kernel<false><<<grid, 32, smem, B.stream>>>(a);
  ALNR_LAUNCHED();
  band_finish_kernel<<<1, 1, 0, B.stream>>>(a, stateful ? 2 : (local ? 1 : 0), fo, c.wave_pair);
  ALNR_LAUNCHED();
#ifdef ALNR_TILE_DEBUG
  {  // per-band time stamps (ns since the first band's entry): entry, woken, first chunk, step 32, step 64, end
    std::vector<unsigned long long> ts((size_t)a.nbands * 8);
    cudaStreamSynchronize(B.stream);
    cudaMemcpy(ts.data(), a.prog, ts.size() * 8, cudaMemcpyDeviceToHost);
    const unsigned long long t0 = ts[0];
    for (uint32_t b = 0; b < a.nbands; b += (a.nbands > 40 ? a.nbands / 20 : 1)) {
      fprintf(stderr, "band %4u:", b);
      for (int k = 0; k < 6; ++k) fprintf(stderr, " %10.1f", (double)(ts[(size_t)b * 8 + k] - t0) / 1000.0);
      if (b) fprintf(stderr, "  | first chunk after the band above: %.1f us", (double)(ts[(size_t)b * 8 + 2] - ts[(size_t)(b - 1) * 8 + 2]) / 1000.0);
      fprintf(stderr, "\n");
    }
  }
#endif
  return ALNR_OK;
}

// Difference-recurrence kernel (kernels_diff.cuh): global mode, DNA-sized alphabet, small integer
// scores — the BASELINE config-4 shape.  Everything else stays on the tile pipeline.
// ALNR_NO_DIFF=1 switches it off (A/B runs, tests).
static bool diff_eligible(const alnr_batch_impl &B, const Chunk &c) {
  if (std::getenv("ALNR_NO_DIFF")) return false;
  if (B.mode != ALNR_GLOBAL || B.V > 4 || c.ckpt_group) return false;
  if (c.m_max < 128 || c.n_max < 128) return false;
  const int del = (int)B.del, ext = (int)B.ext;
  if (!(ext >= 1 && del >= ext && del <= 24)) return false;
  for (double v : B.matrix)
    if (v + 2 * del + ext > 127.0 || v < -1024.0) return false;  // S = s + B + p is a byte
  return true;
}

static int diff_run_pair(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const FillOut &fo) {
  DiffArgs a;
  a.q = B.d_q.p + c.q0; a.t = B.d_t.p + c.t0;
  a.n = c.n_max; a.m = c.m_max;
  a.del = (int)B.del; a.ext = (int)B.ext;
  const int bias = 2 * a.del + a.ext;
  for (int ts = 0; ts < 4; ++ts) {  // one word per TARGET symbol, one byte per query symbol
    uint32_t w = 0;
    for (int qs = 0; qs < 4; ++qs) {
      int v = (qs < B.V && ts < B.V) ? (int)B.matrix[(size_t)ts * B.V + qs] + bias : 0;
      v = std::min(std::max(v, 0), 127);  // a negative S never wins the max against DH, DV >= 0
      w |= (uint32_t)v << (8 * qs);
    }
    a.wtab[ts] = w;
  }
  a.nbands = (a.m + kDiffBandRows - 1) / kDiffBandRows;
  a.npad = (a.n + 63) / 64 * 64;
  // scratch: [pad 2][corner acc][pad to 128 bytes][rows: nbands * npad 16-bit values][256 words of slack for the
  // kernel's reads past the end of a row]
  const size_t words = 32 + ((size_t)a.nbands * a.npad + 1) / 2 + 256;
  if (B.d_band.n < words) {
    if (B.d_band.p) ALNR_CUDA(cudaStreamSynchronize(B.stream));  // an earlier pair may still use the old rows
    ALNR_CUDA(B.d_band.alloc(words));
  }
  a.corner_acc = B.d_band.p + 2;
  a.rows = reinterpret_cast<uint16_t *>(B.d_band.p + 32);  // 128-byte aligned: a pair of groups is one 32-byte sector
  a.stamps = nullptr;
#ifdef ALNR_DIFF_DEBUG
  static unsigned long long *d_stamps = nullptr;
  if (!d_stamps) cudaMalloc(&d_stamps, (size_t)65536 * 64);
  a.stamps = d_stamps;
#endif
  a.D = B.Dp();
  a.dcols = dl.ncols;
  static unsigned resident = 0;
  if (!resident) {
    int occ = 0, sms = 0, dev = 0;
    ALNR_CUDA(cudaGetDevice(&dev));
    ALNR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fill_diff_kernel, 32, 0));
    resident = (unsigned)std::max(1, occ) * (unsigned)std::max(1, sms);
  }
  ALNR_CUDA(cudaMemsetAsync(B.d_band.p, 0, 16, B.stream));
  if (a.nbands > 1)  // every boundary value below row 0 starts as "not produced yet"
    ALNR_CUDA(cudaMemsetAsync(a.rows + a.npad, 0xff, (size_t)(a.nbands - 1) * a.npad * sizeof(uint16_t), B.stream));
  unsigned grid = std::min<unsigned>(a.nbands, resident);  // a waiter depends on lower bands only
  if (const char *e = std::getenv("ALNR_DIFF_GRID")) grid = std::min<unsigned>(grid, (unsigned)std::max(1, std::atoi(e)));
  fill_diff_kernel<<<grid, 32, 0, B.stream>>>(a);
  ALNR_LAUNCHED();
  diff_finish_kernel<<<1, 1, 0, B.stream>>>(a, fo, c.wave_pair);
  ALNR_LAUNCHED();
#ifdef ALNR_DIFF_DEBUG
  {  // per-band time stamps, us since the first band's entry: entry, woken, block 1, t0 = 64, 72, 8192, end
    std::vector<unsigned long long> ts((size_t)a.nbands * 8);
    cudaStreamSynchronize(B.stream);
    cudaMemcpy(ts.data(), a.stamps, ts.size() * 8, cudaMemcpyDeviceToHost);
    const unsigned long long t0 = ts[0];
    for (uint32_t b = 0; b < a.nbands; b += (a.nbands > 40 ? a.nbands / 24 : 1)) {
      fprintf(stderr, "band %4u:", b);
      for (int k = 0; k < 7; ++k) fprintf(stderr, " %9.1f", (double)(ts[(size_t)b * 8 + k] - t0) / 1000.0);
      if (b) fprintf(stderr, " | behind the band above at block 1 / 72 / 8192: %.2f %.2f %.2f us; own block 1 -> 72: %.2f us; above's 72 -> own block 1: %.2f us",
                     (double)(ts[(size_t)b * 8 + 2] - ts[(size_t)(b - 1) * 8 + 2]) / 1000.0,
                     (double)(ts[(size_t)b * 8 + 4] - ts[(size_t)(b - 1) * 8 + 4]) / 1000.0,
                     (double)(ts[(size_t)b * 8 + 5] - ts[(size_t)(b - 1) * 8 + 5]) / 1000.0,
                     (double)(ts[(size_t)b * 8 + 4] - ts[(size_t)b * 8 + 2]) / 1000.0,
                     (double)((long long)ts[(size_t)b * 8 + 2] - (long long)ts[(size_t)(b - 1) * 8 + 4]) / 1000.0);
      fprintf(stderr, "\n");
    }
  }
#endif
  return ALNR_OK;
}

// rows per lane of the tile pipeline (its D layout is DLayout::make_blocked(4, ...))
uint32_t wave_band_rows(uint32_t m) {
  (void)m;
  return 4u;  // measured: 128-row bands beat 64-row bands at 3 x 10^4 and at 10^5 (pipeline fill dominates)
}

int wave_run_pair(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const FillOut &fo) {
  if (diff_eligible(B, c)) return diff_run_pair(B, c, dl, fo);
  return tile_run_pair(B, c, dl, fo);
}

}  // namespace alnr
