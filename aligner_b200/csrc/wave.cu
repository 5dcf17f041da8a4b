// wave.cu — host side of the intra-pair wavefront path for one long pair: the band pipeline
// kernel (kernels_band.cuh).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "kernels_tile.cuh"
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
  if (B.mode == ALNR_LOCAL && B.del != B.ext) return false;  // stateful penalty: columns serialise
  double mag = 0;
  for (double v : B.matrix) mag = std::max(mag, v);
  if (B.mode == ALNR_LOCAL && mag * std::min(m, n) >= 16777216.0) return false;  // 24-bit value field
  return true;
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
  const size_t words = 4 + row_words + dbg_words;
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
  a.D = B.Dp();
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
  if (local) fill_tile_kernel<true><<<grid, 32, smem, B.stream>>>(a);
  else fill_tile_kernel<false><<<grid, 32, smem, B.stream>>>(a);
  ALNR_LAUNCHED();
  band_finish_kernel<<<1, 1, 0, B.stream>>>(a, local ? 1 : 0, fo, c.wave_pair);
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

// Band pipeline (kernels_band.cuh): one persistent launch for the whole pair.
static int band_run_pair(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const FillOut &fo) {
  BandArgs a;
  a.q = B.d_q.p + c.q0; a.t = B.d_t.p + c.t0;
  a.n = c.n_max; a.m = c.m_max; a.V = B.V; a.matrix = B.d_mati.p;
  a.del = (int)B.del; a.ext = (int)B.ext;
  const uint32_t R = dl.rows_per_word, band_rows = 32 * R;
  a.nbands = (a.m + band_rows - 1) / band_rows;
  const bool local = B.mode == ALNR_LOCAL;
  // scratch: [best (8 B)][corner][pad][prog nbands][rows nbands * (n + 1)]
  const size_t words = 4 + (size_t)a.nbands + (size_t)a.nbands * (a.n + 1);
  if (B.d_band.n < words) {
    // a longer pair than the previous one of this batch: its fill may still be running on the old rows
    if (B.d_band.p) ALNR_CUDA(cudaStreamSynchronize(B.stream));
    ALNR_CUDA(B.d_band.alloc(words));
  }
  int *base = B.d_band.p;
  a.best = reinterpret_cast<unsigned long long *>(base);
  a.corner_out = base + 2;
  a.prog = reinterpret_cast<uint32_t *>(base + 4);
  a.rows = base + 4 + a.nbands;
  a.D = B.Dp();
  a.dcols = dl.ncols;
  static unsigned resident = 0;
  if (!resident) {
    int occ0 = 0, occ1 = 0, sms = 0, dev = 0;
    ALNR_CUDA(cudaGetDevice(&dev));
    ALNR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const size_t smem0 = 32 * 32 * 4;
    ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ0, fill_band_kernel<4, false>, 32, smem0));
    ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, fill_band_kernel<4, true>, 32, smem0));
    resident = (unsigned)std::max(1, std::min(occ0, occ1)) * (unsigned)std::max(1, sms);
  }
  band_init_kernel<<<(a.nbands + 255) / 256, 256, 0, B.stream>>>(a);
  ALNR_LAUNCHED();
  const unsigned grid = std::min<unsigned>(a.nbands, resident);  // every block must be resident
  const size_t smem = (size_t)B.V * B.V * 4;
  if (R == 2) {
    if (local) fill_band_kernel<2, true><<<grid, 32, smem, B.stream>>>(a);
    else fill_band_kernel<2, false><<<grid, 32, smem, B.stream>>>(a);
  } else {
    if (local) fill_band_kernel<4, true><<<grid, 32, smem, B.stream>>>(a);
    else fill_band_kernel<4, false><<<grid, 32, smem, B.stream>>>(a);
  }
  ALNR_LAUNCHED();
  band_finish_kernel<<<1, 1, 0, B.stream>>>(a, local ? 1 : 0, fo, c.wave_pair);
  ALNR_LAUNCHED();
  return ALNR_OK;
}

// rows per lane of the band pipeline: thinner bands = more warps in flight (ALNR_BAND_R overrides)
uint32_t wave_band_rows(uint32_t m) {
  static const int env = std::getenv("ALNR_BAND_R") ? std::atoi(std::getenv("ALNR_BAND_R")) : 0;
  if (env == 2) return 2u;  // first-generation column kernel with 64-row bands
  (void)m;
  return 4u;  // measured: 128-row bands beat 64-row bands at 3 x 10^4 and at 10^5 (pipeline fill dominates)
}

int wave_run_pair(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const FillOut &fo) {
  // ALNR_BAND_COLUMN=1 / ALNR_BAND_R=2: the first-generation column-per-step kernel (A/B runs)
  static const bool column = std::getenv("ALNR_BAND_COLUMN") != nullptr;
  if (column || dl.rows_per_word != 4) return band_run_pair(B, c, dl, fo);
  return tile_run_pair(B, c, dl, fo);
}

}  // namespace alnr
