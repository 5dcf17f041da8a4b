// wave.cu — host side of the intra-pair wavefront kernel (kernels_wave.cuh).
#include <algorithm>
#include <cstdlib>

#include "kernels_wave.cuh"
#include "strip_launch.h"
#include "wave_launch.h"

namespace alnr {

static const uint64_t kWaveMinCells = 1ull << 24;  // below this the per-thread kernels win

bool wave_eligible(const alnr_batch_impl &B, uint32_t m, uint32_t n) {
  if (B.numeric != ALNR_NUM_I32 || B.want_matrices) return false;
  if ((uint64_t)m * n < kWaveMinCells) return false;
  if (m >= 0xFFFFFu || n >= 0xFFFFFu) return false;       // arg-max key packs 20-bit coordinates
  if (B.mode == ALNR_LOCAL && B.del != B.ext) return false;  // stateful penalty: columns serialise
  double mag = 0;
  for (double v : B.matrix) mag = std::max(mag, v);
  if (B.mode == ALNR_LOCAL && mag * std::min(m, n) >= 16777216.0) return false;  // 24-bit value field
  return true;
}

int wave_run_pair(alnr_batch_impl &B, const Chunk &c, const DLayout &dl, const FillOut &fo) {
  static const int R = std::getenv("ALNR_WAVE_R8") ? 8 : 4;
  static const uint32_t TWenv = std::getenv("ALNR_WAVE_TW") ? (uint32_t)std::atoi(std::getenv("ALNR_WAVE_TW")) : 128;
  const uint32_t TH = 32 * R, TW = std::max(32u, TWenv / 32 * 32);
  WaveArgs a;
  a.q = B.d_q.p + c.q0; a.t = B.d_t.p + c.t0;
  a.n = c.n_max; a.m = c.m_max; a.V = B.V; a.matrix = B.d_mati.p;
  a.del = (int)B.del; a.ext = (int)B.ext; a.local = B.mode == ALNR_LOCAL;
  a.TW = TW;
  a.tile_rows = (a.m + TH - 1) / TH; a.tile_cols = (a.n + TW - 1) / TW;
  int *base = B.hcol_i();
  a.rowbuf = base; a.colbuf = base + (a.n + 1); a.corner = a.colbuf + (a.m + 1);
  a.corner_out = a.corner + a.tile_rows;
  a.best = reinterpret_cast<unsigned long long *>(base + (((size_t)(a.n + 1) + (a.m + 1) + a.tile_rows + 1 + 1) / 2 * 2));
  a.Dbytes = reinterpret_cast<uint8_t *>(B.Dp());
  a.ncols = dl.ncols;
  a.diag = 0;
  const uint32_t mx = std::max(std::max(a.n, a.m) + 1, a.tile_rows);
  wave_init_kernel<<<(mx + 255) / 256, 256, 0, B.stream>>>(a, TH);
  ALNR_LAUNCHED();
  const size_t smem = (size_t)B.V * B.V * 4 + 4 * (size_t)TW;
  for (uint32_t d = 0; d < a.tile_rows + a.tile_cols - 1; ++d) {
    a.diag = d;
    const uint32_t i_lo = d >= a.tile_cols ? d - a.tile_cols + 1 : 0, i_hi = std::min(d, a.tile_rows - 1);
    const uint32_t tiles = i_hi - i_lo + 1;
    const unsigned grid = (tiles + 3) / 4;
    if (a.local) {
      if (R == 8) fill_wave_kernel<8, true><<<grid, 128, smem, B.stream>>>(a);
      else fill_wave_kernel<4, true><<<grid, 128, smem, B.stream>>>(a);
    } else {
      if (R == 8) fill_wave_kernel<8, false><<<grid, 128, smem, B.stream>>>(a);
      else fill_wave_kernel<4, false><<<grid, 128, smem, B.stream>>>(a);
    }
    ALNR_LAUNCHED();
  }
  wave_finish_kernel<<<1, 1, 0, B.stream>>>(a, fo, c.wave_pair);
  ALNR_LAUNCHED();
  return ALNR_OK;
}

}  // namespace alnr
