// heur.cu — the heuristic matrix-search loop on the GPU:
//   HeuristicAligner::perform_alignment   aligner-core/src/heuristic/mod.rs:36-78
//   outer search over candidate matrices  aligner-web/src/bin/dispatcher/handlers.rs:193-251,
//                                         aligner-web/src/bin/node/main.rs:61-74
// One ITEM = (sequence pair, candidate matrix).  Per round every active item runs one FP64
// local alignment with its own current matrix (fill_exact_kernel<double>, per-item matrices),
// the traceback accumulates the item's V x V frequency matrix (alignment.rs:13-23), and an
// update kernel either stops the item (f did not improve: heuristic/mod.rs:73-76) or derives its
// next matrix with transform_matrix (mod.rs:64-72) and keeps it on the active list.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <vector>

#include "host_util.h"
#include "kernels_exact.cuh"
#include "kernels_trace.cuh"
#include "transform.cuh"
#include "warp_args.h"

namespace alnr {

struct HeurState {
  const uint32_t *pair_idx, *mat_idx;  // per item
  const double *matrices;              // K * VV candidates
  const double *freqs;                 // pairs * V
  double kd, r2;
  int V;
  double *Tcur;      // items * VV: matrix of the NEXT / current pass
  double *Tlast;     // items * VV: matrix that produced the pass being judged
  double *freqm;     // items * VV: frequency matrix of the current pass
  double *maxf, *fcur, *fout;
  int32_t *status, *passes;
  uint64_t *coords_cur, *coords_out;
  const int32_t *pair_status;
  // Items are grouped by the rows-per-lane instantiation their target needs (pair_gid[pair]); every
  // group keeps its own active list (list base + gid * list_stride, counter base + gid).
  const uint8_t *pair_gid;
  uint32_t list_stride;
};

constexpr int kHeurGroups = 4;

// Both kernels below give every item a group of 8 lanes (transform_matrix_g8): grid = items * 8 threads.
static __global__ void heur_init_kernel(HeurState h, uint32_t items, uint32_t *list, uint32_t *count) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t i = tid >> 3;
  if (i >= items) return;
  const int gl = threadIdx.x & 7;
  const unsigned gmask = 0xffu << (threadIdx.x & 24);
  const int VV = h.V * h.V;
  int st = h.pair_status[h.pair_idx[i]];
  if (!st) st = transform_matrix_g8(h.matrices + (size_t)h.mat_idx[i] * VV, h.V, h.V, h.kd, h.r2,
                                    h.freqs + (size_t)h.pair_idx[i] * h.V, h.Tcur + (size_t)i * VV, gmask, gl);
  if (gl != 0) return;
  h.status[i] = st;
  h.maxf[i] = 0.0;  // heuristic/mod.rs:55
  h.passes[i] = 0;
  h.fout[i] = 0.0;
  const uint32_t g = h.pair_gid[h.pair_idx[i]];
  if (!st) list[(size_t)g * h.list_stride + atomicAdd(count + g, 1u)] = i;
}

// after fill + traceback of one round
static __global__ void heur_update_kernel(HeurState h, const uint32_t *list, uint32_t n_active,
                                          uint32_t *next, uint32_t *next_count) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t k = tid >> 3;
  if (k >= n_active) return;
  const int gl = threadIdx.x & 7;
  const unsigned gmask = 0xffu << (threadIdx.x & 24);
  const uint32_t i = list[k];
  const int VV = h.V * h.V;
  if (h.status[i] != 0) return;  // e.g. 11: no positive cell — the reference panics (simple/mod.rs:214)
  const double f = h.fcur[i];
  const bool improved = f > h.maxf[i];  // heuristic/mod.rs:64
  __syncwarp(gmask);                    // every lane has read maxf before lane 0 updates it
  double *T = h.Tcur + (size_t)i * VV, *L = h.Tlast + (size_t)i * VV;
  if (improved) {
    const int st = transform_matrix_g8(h.freqm + (size_t)i * VV, h.V, h.V, h.kd, h.r2,
                                       h.freqs + (size_t)h.pair_idx[i] * h.V, T, gmask, gl);  // :66-72
    if (gl != 0) return;
    h.passes[i] += 1;
    h.maxf[i] = f;
    if (st) { h.status[i] = st; return; }  // .unwrap() panic in the reference
    const uint32_t g = h.pair_gid[h.pair_idx[i]];
    next[(size_t)g * h.list_stride + atomicAdd(next_count + g, 1u)] = i;
  } else {  // :73-76 — the LAST (non-improving) result and the matrix that produced it
    for (int j = gl; j < VV; j += 8) L[j] = T[j];
    if (gl != 0) return;
    h.passes[i] += 1;
    h.fout[i] = f;
    for (int j = 0; j < 4; ++j) h.coords_out[4 * (size_t)i + j] = h.coords_cur[4 * (size_t)i + j];
  }
}

// items still improving when the round cap is reached: per-item status, the last pass as result
static __global__ void heur_mark_unconverged_kernel(HeurState h, const uint32_t *list, uint32_t n_active) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_active) return;
  const uint32_t i = list[k];
  if (h.status[i] != 0) return;
  h.status[i] = ALNR_CALCULATION_ERROR;
  h.fout[i] = h.maxf[i];
}

}  // namespace alnr

using namespace alnr;

extern "C" int alnr_transform_matrix(const double *matrix, int rows, int cols, double kd,
                                     double r_squared, const double *frequencies, double *out) {
  if (!matrix || !frequencies || !out || rows < 1 || cols < 1) return ALNR_INVALID_ARGUMENT;
  // host instance of the same routine the kernels run (aligner-helpers/src/matrices/mod.rs:19-68)
  std::vector<double> tmp((size_t)rows * cols);
  int st = transform_matrix_rc(matrix, rows, cols, kd, r_squared, frequencies, tmp.data());
  if (st) return ALNR_WRONG_MATRIX;
  std::copy(tmp.begin(), tmp.end(), out);
  return ALNR_OK;
}

extern "C" int alnr_heuristic_batch(const alnr_params *p, const alnr_batch *b,
                                    const alnr_heuristic_items *it, alnr_heuristic_out *out) {
  if (!p || !b || !it || !out || !it->matrices || !it->frequencies) return ALNR_INVALID_ARGUMENT;
  if (b->count && (!b->q || !b->t || !b->q_off || !b->t_off)) return ALNR_INVALID_ARGUMENT;  // byte codes only
  if (p->volume < 1 || p->volume > 32) return ALNR_MATRIX_SHAPE_ERROR;
  if (!(std::fabs(p->del) < 1e290 && std::fabs(p->ext) < 1e290)) return ALNR_INVALID_ARGUMENT;  // NaN / inf penalties
  const int V = p->volume, VV = V * V;
  const uint64_t P = b->count, K = it->n_matrices;
  const bool grid = it->pair_idx == nullptr;  // full grid: item = pair * K + k
  const uint64_t items = grid ? P * K : it->count;
  if (items == 0) return ALNR_OK;
  if (items >= (1ull << 31) || P >= (1ull << 31)) return ALNR_INVALID_ARGUMENT;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    (void)cudaGetLastError();
    set_error_detail("no usable CUDA device (this library has no CPU fallback)");
    return ALNR_CUDA_ERROR;
  }
  const int device = p->device < 0 ? 0 : p->device;
  if (device >= ndev) return ALNR_INVALID_ARGUMENT;
  ALNR_CUDA(cudaSetDevice(device));
  double r2 = it->r_squared;
  if (std::fabs(r2 - 0.0) < DBL_EPSILON) r2 = (double)VV;  // heuristic/mod.rs:48-50

  uint32_t n_max = 0, m_max = 0;
  for (uint64_t i = 0; i < P; ++i) {
    n_max = std::max<uint32_t>(n_max, (uint32_t)(b->q_off[i + 1] - b->q_off[i]));
    m_max = std::max<uint32_t>(m_max, (uint32_t)(b->t_off[i + 1] - b->t_off[i]));
  }
  // Groups by rows per lane: a 300-row target in the 352-row band of the longest target (R = 11)
  // drags 15 % idle rows through every step; with its own R = 10 instantiation it drags 6 %.  At most
  // kHeurGroups groups (the rarest R values are merged upwards); all groups share the rounds — one
  // launch sequence per group and round, one host synchronisation per round.  ALNR_HEUR_ONE_GROUP=1:
  // one group with the R of the longest target (the round-1 behaviour, A/B runs).
  std::vector<uint8_t> pair_gid(P, 0);
  uint32_t group_R[kHeurGroups] = {0, 0, 0, 0}, group_m[kHeurGroups] = {0, 0, 0, 0};
  int n_groups = 1;
  {
    uint64_t hist[17] = {0};
    for (uint64_t i = 0; i < P; ++i) hist[warp_pick_rows(std::max<uint32_t>((uint32_t)(b->t_off[i + 1] - b->t_off[i]), 1u))]++;
    int up[17];  // R -> the R of the group it is merged into
    for (int r = 0; r <= 16; ++r) up[r] = r;
    auto distinct = [&] { int c = 0; for (int r = 2; r <= 16; ++r) c += hist[r] != 0; return c; };
    const bool one_group = getenv("ALNR_HEUR_ONE_GROUP") != nullptr;
    while (distinct() > (one_group ? 1 : kHeurGroups)) {
      int best = -1;  // the rarest R that has a larger neighbour
      for (int r = 2; r < 16; ++r) {
        if (!hist[r]) continue;
        bool has_up = false;
        for (int q = r + 1; q <= 16; ++q) has_up = has_up || hist[q];
        if (has_up && (best < 0 || hist[r] < hist[best])) best = r;
      }
      int nxt = best + 1;
      while (!hist[nxt]) ++nxt;
      hist[nxt] += hist[best]; hist[best] = 0;
      for (int r = 2; r <= 16; ++r) if (up[r] == best) up[r] = nxt;
    }
    int gid_of[17];
    n_groups = 0;
    for (int r = 2; r <= 16; ++r) if (hist[r]) { gid_of[r] = n_groups; group_R[n_groups++] = (uint32_t)r; }
    for (uint64_t i = 0; i < P; ++i) {
      const uint32_t m = std::max<uint32_t>((uint32_t)(b->t_off[i + 1] - b->t_off[i]), 1u);
      const int g = gid_of[up[warp_pick_rows(m)]];
      pair_gid[i] = (uint8_t)g;
      group_m[g] = std::max(group_m[g], m);
    }
  }
  const uint64_t q_total = b->q_off[P], t_total = b->t_off[P];
  // one warp per item (kernels_warp.cuh): the item's own matrix sits in shared memory
  const DLayout dl0 = DLayout::make_skewed(warp_pick_rows(std::max(m_max, 1u)), std::max(m_max, 1u), std::max(n_max, 1u));
  const bool stateful = p->del != p->ext;
  const uint32_t nzw = (n_max + 31) / 32;
  const uint64_t row_words = (uint64_t)std::max(m_max, n_max) + 1;  // band rows / exact-kernel column
  const uint64_t per_item = dl0.words_for(1) * 4 + 8ull * row_words + 3ull * VV * 8 + 132 + (stateful ? 8ull * nzw : 0);
  size_t free_b = 0, total_b = 0;
  ALNR_CUDA(cudaMemGetInfo(&free_b, &total_b));
  const uint64_t budget = std::max<uint64_t>(std::min<uint64_t>(free_b / 3, 16ull << 30), 64ull << 20);
  uint64_t chunk = std::max<uint64_t>(budget / per_item / 32 * 32, 32);
  chunk = std::min<uint64_t>(chunk, std::min<uint64_t>((items + 31) / 32 * 32, 1u << 20));

  cudaStream_t stream;
  ALNR_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamSynchronize(s); cudaStreamDestroy(s); } } guard{stream};

  DevBuf<uint8_t> d_q, d_t, d_gid;
  DevBuf<uint64_t> d_qoff, d_toff, d_cc, d_co;
  DevBuf<uint32_t> d_pi, d_mi, d_lists, d_cnt, d_start, d_D, d_wz;
  DevBuf<int32_t> d_pst, d_st, d_passes;
  DevBuf<int> d_flag;
  DevBuf<double> d_mats, d_freqs, d_T, d_L, d_F, d_maxf, d_fcur, d_fout, d_corner, d_hcol;
  // declared after the buffers, so destroyed first: nothing returns to the pool while kernels of an
  // early return are still queued
  struct SyncGuard { cudaStream_t s; ~SyncGuard() { cudaStreamSynchronize(s); (void)cudaGetLastError(); } } sync_guard{stream};
  ALNR_CUDA(d_q.alloc(q_total + 16)); ALNR_CUDA(d_t.alloc(t_total + 16));
  ALNR_CUDA(d_qoff.alloc(P + 1)); ALNR_CUDA(d_toff.alloc(P + 1));
  ALNR_CUDA(d_mats.alloc(K * VV)); ALNR_CUDA(d_freqs.alloc(P * V));
  ALNR_CUDA(d_pst.alloc(P)); ALNR_CUDA(d_flag.alloc(4));
  ALNR_CUDA(d_pi.alloc(chunk)); ALNR_CUDA(d_mi.alloc(chunk));
  ALNR_CUDA(d_lists.alloc(2 * chunk * kHeurGroups)); ALNR_CUDA(d_cnt.alloc(64 * kHeurGroups));
  ALNR_CUDA(d_gid.alloc(P));
  ALNR_CUDA(d_start.alloc(2 * chunk)); ALNR_CUDA(d_D.alloc(dl0.words_for(chunk)));
  ALNR_CUDA(d_st.alloc(chunk)); ALNR_CUDA(d_passes.alloc(chunk));
  ALNR_CUDA(d_T.alloc(chunk * VV)); ALNR_CUDA(d_L.alloc(chunk * VV)); ALNR_CUDA(d_F.alloc(chunk * VV));
  ALNR_CUDA(d_maxf.alloc(chunk)); ALNR_CUDA(d_fcur.alloc(chunk)); ALNR_CUDA(d_fout.alloc(chunk));
  ALNR_CUDA(d_corner.alloc(chunk)); ALNR_CUDA(d_hcol.alloc((chunk + 31) / 32 * 32 * row_words));
  // fail list: items whose speculation never settles or whose matrix is not finite (exact kernel)
  ALNR_CUDA(d_wz.alloc(warp_scratch_words(chunk, nzw, stateful)));
  ALNR_CUDA(d_cc.alloc(chunk * 4)); ALNR_CUDA(d_co.alloc(chunk * 4));

  auto h2d = [&](void *dst, const void *src, size_t bytes) {
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream);
  };
  ALNR_CUDA(h2d(d_q.p, b->q, q_total)); ALNR_CUDA(h2d(d_t.p, b->t, t_total));
  ALNR_CUDA(h2d(d_qoff.p, b->q_off, (P + 1) * 8)); ALNR_CUDA(h2d(d_toff.p, b->t_off, (P + 1) * 8));
  ALNR_CUDA(h2d(d_mats.p, it->matrices, K * VV * 8)); ALNR_CUDA(h2d(d_freqs.p, it->frequencies, P * V * 8));
  ALNR_CUDA(h2d(d_gid.p, pair_gid.data(), P));
  // per-pair validation (empty sequences, symbols outside the alphabet)
  ALNR_CUDA(cudaMemsetAsync(d_flag.p, 0, 4 * sizeof(int), stream));
  any_symbol_out_of_range_kernel<<<148 * 4, 256, 0, stream>>>(d_q.p, q_total, V, d_flag.p); ALNR_LAUNCHED();
  any_symbol_out_of_range_kernel<<<148 * 4, 256, 0, stream>>>(d_t.p, t_total, V, d_flag.p); ALNR_LAUNCHED();
  {
    ChunkView cvp{d_q.p, d_t.p, d_qoff.p, d_toff.p, nullptr, 0, (uint32_t)P};
    validate_kernel<<<(unsigned)((P + 127) / 128), 128, 0, stream>>>(cvp, V, d_pst.p, d_flag.p); ALNR_LAUNCHED();
  }

  std::vector<uint32_t> hpi(chunk), hmi(chunk);
  for (uint64_t base = 0; base < items; base += chunk) {
    const uint32_t cnt = (uint32_t)std::min<uint64_t>(chunk, items - base);
    for (uint32_t i = 0; i < cnt; ++i) {
      const uint64_t g = base + i;
      hpi[i] = grid ? (uint32_t)(g / K) : it->pair_idx[g];
      hmi[i] = grid ? (uint32_t)(g % K) : it->mat_idx[g];
      if (hpi[i] >= P || hmi[i] >= K) return ALNR_INVALID_ARGUMENT;
    }
    ALNR_CUDA(h2d(d_pi.p, hpi.data(), cnt * 4)); ALNR_CUDA(h2d(d_mi.p, hmi.data(), cnt * 4));
    ALNR_CUDA(cudaStreamSynchronize(stream));  // hpi/hmi are reused by the next chunk
    HeurState hs{d_pi.p, d_mi.p, d_mats.p, d_freqs.p, it->kd, r2, V, d_T.p, d_L.p, d_F.p, d_maxf.p,
                 d_fcur.p, d_fout.p, d_st.p, d_passes.p, d_cc.p, d_co.p, d_pst.p, d_gid.p, (uint32_t)chunk};
    uint32_t *lists[2] = {d_lists.p, d_lists.p + chunk * kHeurGroups};
    ALNR_CUDA(cudaMemsetAsync(d_cnt.p, 0, 64 * kHeurGroups * sizeof(uint32_t), stream));
    ALNR_CUDA(cudaMemsetAsync(d_co.p, 0, (size_t)cnt * 32, stream));
    heur_init_kernel<<<(cnt + 15) / 16, 128, 0, stream>>>(hs, cnt, lists[0], d_cnt.p); ALNR_LAUNCHED();
    uint32_t act[kHeurGroups] = {0, 0, 0, 0};  // active items per group
    ALNR_CUDA(cudaMemcpyAsync(act, d_cnt.p, sizeof(act), cudaMemcpyDeviceToHost, stream));
    ALNR_CUDA(cudaStreamSynchronize(stream));
    auto total_active = [&] { uint32_t t = 0; for (int g = 0; g < n_groups; ++g) t += act[g]; return t; };
    uint32_t n_active = total_active();
    // The reference loops until f stops improving with no cap (heuristic/mod.rs:58-77); f is bounded
    // above, so it ends, but a cap keeps a pathological item from holding the batch: items still
    // active after max_rounds passes get the per-item status ALNR_CALCULATION_ERROR, every other
    // item keeps its result and the call returns ALNR_OK.  ALNR_HEUR_MAX_ROUNDS (default 60; the
    // round counters limit it to 62).  Deviation from the reference, documented in DESIGN.md.
    static const int max_rounds = [] {
      const char *e = getenv("ALNR_HEUR_MAX_ROUNDS");
      const int v = e ? atoi(e) : 60;
      return v < 1 ? 1 : (v > 62 ? 62 : v);
    }();
    for (int round = 0; n_active > 0; ++round) {
      if (round >= max_rounds) {
        for (int g = 0; g < n_groups; ++g) {
          if (!act[g]) continue;
          heur_mark_unconverged_kernel<<<(act[g] + 127) / 128, 128, 0, stream>>>(hs, lists[round & 1] + (size_t)g * chunk, act[g]);
          ALNR_LAUNCHED();
        }
        break;
      }
      uint32_t *cnt_next = d_cnt.p + (size_t)(1 + round) * kHeurGroups;
      for (int g = 0; g < n_groups; ++g) {
        const uint32_t n_g = act[g];
        if (n_g == 0) continue;
        const uint32_t *cur = lists[round & 1] + (size_t)g * chunk;
        uint32_t *nxt = lists[(round + 1) & 1];
        const uint32_t m_g = std::max(group_m[g], 1u);
        const DLayout dl_g = DLayout::make_skewed(group_R[g], m_g, std::max(n_max, 1u));
        ChunkView cv{d_q.p, d_t.p, d_qoff.p, d_toff.p, cur, 0, n_g};
        cv.seqmap = d_pi.p;
        FillOut fo{d_fcur.p, d_corner.p, d_start.p, d_st.p};
        WarpArgs<double> w;
        w.cv = cv; w.V = V; w.del = p->del; w.ext = p->ext; w.stateful = stateful;
        w.matrix = nullptr; w.item_matrices = d_T.p; w.D = d_D.p; w.dl = dl_g; w.fo = fo;
        w.rowbuf = d_hcol.p; w.rowbuf_stride = n_max + 1;
        w.nzw = nzw;
        warp_scratch_bind(w, d_wz.p, chunk, true);
        w.H_out = nullptr; w.D_out = nullptr;
        ALNR_CUDA(cudaMemsetAsync(d_wz.p, 0, 8 * sizeof(uint32_t), stream));
        { int wst = warp_launch_f64(stream, w, true, false); if (wst) return wst; }
        {  // items the warp kernel handed over (fail list): exact column-major kernel
          ExactArgs<double> e;
          e.cv = cv; e.mode = ALNR_LOCAL; e.V = V; e.del = p->del; e.ext = p->ext;
          e.matrix = nullptr; e.hcol = d_hcol.p; e.hrows = m_max + 1; e.D = d_D.p; e.dl = dl_g; e.fo = fo;
          e.H_out = nullptr; e.D_out = nullptr; e.item_matrices = d_T.p;
          e.list = w.fail_list; e.list_count = w.fail_count;
          fill_exact_kernel<double><<<(n_g + 127) / 128, 128, (size_t)VV * 8, stream>>>(e); ALNR_LAUNCHED();
        }
        TraceOut to{d_cc.p, nullptr, nullptr, nullptr, nullptr, d_F.p};
        TraceArgs ta{cv, ALNR_LOCAL, V, d_D.p, dl_g, fo, to};
        // one warp per item, counts in shared memory (kernels_trace.cuh); very long pairs whose staged
        // sequences do not fit shared memory take the one-thread-per-item kernel
        const size_t tsm = 4 * ((size_t)VV + ((n_max + 7) / 8 * 8 + (m_g + 7) / 8 * 8) / 8) * sizeof(double);
        if (tsm <= 200 * 1024) {
          if (tsm > 48 * 1024)
            ALNR_CUDA(cudaFuncSetAttribute(traceback_freq_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm));
          traceback_freq_warp_kernel<<<(n_g + 3) / 4, 128, tsm, stream>>>(ta, n_max, m_g); ALNR_LAUNCHED();
        } else {
          traceback_kernel<<<(n_g + 127) / 128, 128, 0, stream>>>(ta); ALNR_LAUNCHED();
        }
        heur_update_kernel<<<(n_g + 15) / 16, 128, 0, stream>>>(hs, cur, n_g, nxt, cnt_next); ALNR_LAUNCHED();
      }
      ALNR_CUDA(cudaMemcpyAsync(act, cnt_next, sizeof(act), cudaMemcpyDeviceToHost, stream));
      ALNR_CUDA(cudaStreamSynchronize(stream));
      n_active = total_active();
    }
    auto d2h = [&](void *dst, const void *src, size_t bytes) -> cudaError_t {
      if (!dst) return cudaSuccess;
      return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream);
    };
    ALNR_CUDA(d2h(out->f ? out->f + base : nullptr, d_fout.p, (size_t)cnt * 8));
    ALNR_CUDA(d2h(out->passes ? out->passes + base : nullptr, d_passes.p, (size_t)cnt * 4));
    ALNR_CUDA(d2h(out->status ? out->status + base : nullptr, d_st.p, (size_t)cnt * 4));
    ALNR_CUDA(d2h(out->coords ? out->coords + base * 4 : nullptr, d_co.p, (size_t)cnt * 32));
    ALNR_CUDA(d2h(out->matrices ? out->matrices + base * VV : nullptr, d_L.p, (size_t)cnt * VV * 8));
    ALNR_CUDA(cudaStreamSynchronize(stream));
  }
  return ALNR_OK;
}

// Fitness reduction in a FIXED order so that the f64 sums do not depend on how many GPUs shared
// the pairs: pairs are cut into blocks of ALNR_FITNESS_BLOCK consecutive pairs (global pair
// index), block_sums[b][k] adds f(pair, k) over the block in pair order, and fitness[k] adds the
// blocks in block order.  A shard that starts on a block boundary produces its blocks exactly;
// the ranks exchange them (every block has one owner, so an all-reduce that adds zeros is exact).
extern "C" int alnr_fitness_block_sums(const double *f_grid, const int32_t *status_grid, uint64_t n_pairs,
                                       uint32_t n_matrices, double *block_sums) {
  if (!f_grid || !block_sums || n_matrices == 0) return ALNR_INVALID_ARGUMENT;
  const uint64_t K = n_matrices, nb = (n_pairs + ALNR_FITNESS_BLOCK - 1) / ALNR_FITNESS_BLOCK;
  for (uint64_t b = 0; b < nb; ++b) {
    double *out = block_sums + b * K;
    for (uint64_t k = 0; k < K; ++k) out[k] = 0.0;
    const uint64_t lo = b * ALNR_FITNESS_BLOCK, hi = std::min<uint64_t>(lo + ALNR_FITNESS_BLOCK, n_pairs);
    for (uint64_t i = lo; i < hi; ++i)
      for (uint64_t k = 0; k < K; ++k) {
        if (status_grid && status_grid[i * K + k] != 0) continue;  // items on which the reference panics
        out[k] = out[k] + f_grid[i * K + k];
      }
  }
  return ALNR_OK;
}

extern "C" int alnr_fitness_from_blocks(const double *block_sums, uint64_t n_blocks, uint32_t n_matrices,
                                        double *fitness) {
  if (!block_sums || !fitness) return ALNR_INVALID_ARGUMENT;
  for (uint64_t k = 0; k < n_matrices; ++k) {
    double acc = 0.0;
    for (uint64_t b = 0; b < n_blocks; ++b) acc = acc + block_sums[b * n_matrices + k];
    fitness[k] = acc;
  }
  return ALNR_OK;
}

// Outer search (dispatcher/handlers.rs:193-251 + node/main.rs:61-74 + the arg-best of
// repository/queries.rs:74-80, taken as a MAXIMUM — the reference's ascending sort is a defect):
// every pair under every candidate matrix, fitness[k] = sum over pairs of f(pair, k) in the fixed
// blocked order above, best_k[pair] = first k with the largest f.
extern "C" int alnr_matrix_search(const alnr_params *p, const alnr_batch *b, const double *matrices,
                                  uint32_t n_matrices, double kd, double r_squared,
                                  const double *frequencies, double *fitness, uint32_t *best_k,
                                  double *best_f, double *f_grid, int32_t *status_grid) {
  if (!b || !fitness || n_matrices == 0) return ALNR_INVALID_ARGUMENT;
  const uint64_t P = b->count, K = n_matrices;
  std::vector<double> fbuf;
  std::vector<int32_t> sbuf;
  double *fg = f_grid;
  int32_t *sg = status_grid;
  if (!fg) { fbuf.resize(P * K); fg = fbuf.data(); }
  if (!sg) { sbuf.resize(P * K); sg = sbuf.data(); }
  alnr_heuristic_items it{nullptr, nullptr, 0, matrices, n_matrices, kd, r_squared, frequencies};
  alnr_heuristic_out out{fg, nullptr, sg, nullptr, nullptr};
  int st = alnr_heuristic_batch(p, b, &it, &out);
  if (st) return st;
  // fitness in the fixed blocked order (see alnr_fitness_block_sums): identical for 1 or N GPUs
  const uint64_t nb = (P + ALNR_FITNESS_BLOCK - 1) / ALNR_FITNESS_BLOCK;
  std::vector<double> bs(nb * K);
  st = alnr_fitness_block_sums(fg, sg, P, n_matrices, bs.data());
  if (st) return st;
  st = alnr_fitness_from_blocks(bs.data(), nb, n_matrices, fitness);
  if (st) return st;
  for (uint64_t i = 0; i < P; ++i) {
    uint32_t bk = 0;
    double bf = -INFINITY;
    for (uint64_t k = 0; k < K; ++k) {
      if (sg[i * K + k] != 0) continue;
      const double f = fg[i * K + k];
      if (f > bf) { bf = f; bk = (uint32_t)k; }
    }
    if (best_k) best_k[i] = bk;
    if (best_f) best_f[i] = bf;
  }
  return ALNR_OK;
}
