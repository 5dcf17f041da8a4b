// warp_args.h — argument block and host entry points of the warp-per-pair fill kernel
// (kernels_warp.cuh; instantiated in warp_i32.cu and warp_f64.cu).
#pragma once
#include "common.cuh"

namespace alnr {

constexpr int kWarpsPerBlock = 4;
constexpr int kWarpPasses = 6;

template <typename T>
struct WarpArgs {
  ChunkView cv;
  int V;
  T del, ext;
  int stateful;            // local mode with del != ext
  const T *matrix;         // V*V shared by all items (device) ...
  const T *item_matrices;  // ... or one V*V matrix per result index (heuristic search)
  uint32_t *D;
  DLayout dl;              // skewed, rows_per_word == R
  FillOut fo;
  T *rowbuf;               // band boundary rows, [slot * rowbuf_stride + x]
  uint32_t rowbuf_stride;
  uint32_t *zbuf;          // [slot][2][nzw] assumed / observed bits (stateful only)
  uint32_t nzw;
  uint32_t *fail_list;     // slots whose speculation did not settle (-> exact kernel)
  uint32_t *fail_count;
  // Incremental speculation (stateful, one band, more rows than one lane holds; see
  // warp_verify_kernel): the first launch makes ONE pass, keeps lane 0's bottom row in rowbuf and
  // puts mispredicted slots on vlist; the verify kernel settles most of them and puts the rest on
  // rlist; a second launch (list = rlist) loops as before.  All nullptr: the warp loops by itself.
  uint32_t *vlist = nullptr, *vcount = nullptr;
  uint32_t *rlist = nullptr, *rcount = nullptr;
  const uint32_t *list = nullptr, *list_count = nullptr;  // second launch: slots to process
  double *H_out;           // optional, single-pair calls: (m+1)*(n+1) row-major
  uint8_t *D_out;
};

// Rows per lane for targets of up to m_max rows: the smallest R in 2..16 with 32 * R >= m_max
// (one band, as few idle rows as possible: 330-row targets take R = 11, 352 rows), else 16
// (several bands).
inline uint32_t warp_pick_rows(uint32_t m_max) {
  const uint32_t r = (m_max + 31) / 32;
  return r < 2 ? 2 : (r > 16 ? 16 : r);
}

// Layout of the speculation scratch: [8 counters][fail list][verify list][re-run list][z bits 2*nzw per slot]
inline uint64_t warp_scratch_words(uint64_t slots, uint32_t nzw, bool stateful) {
  return 8 + (stateful ? 3 * slots + 2ull * nzw * slots : slots);
}
template <typename T>
inline void warp_scratch_bind(WarpArgs<T> &a, uint32_t *wz, uint64_t slots, bool incremental) {
  a.fail_count = wz; a.fail_list = wz ? wz + 8 : nullptr;
  a.zbuf = (wz && a.stateful) ? wz + 8 + 3 * slots : nullptr;
  const bool inc = wz && a.stateful && incremental;
  a.vcount = inc ? wz + 1 : nullptr; a.vlist = inc ? wz + 8 + slots : nullptr;
  a.rcount = inc ? wz + 2 : nullptr; a.rlist = inc ? wz + 8 + 2 * slots : nullptr;
}

// Enqueue the warp-per-pair fill of a.cv.count slots on `stream`; a.dl.rows_per_word selects R.
// mats: also write a.H_out / a.D_out (single pair; R must be 8).
int warp_launch_i32(cudaStream_t stream, const WarpArgs<int> &a, bool local, bool mats);
int warp_launch_f64(cudaStream_t stream, const WarpArgs<double> &a, bool local, bool mats);

}  // namespace alnr
