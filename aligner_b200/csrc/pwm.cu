// pwm.cu — PWM aligner over windows of one HBM-resident sequence (include/aligner_b200.h, "PWM
// aligner"; reference: aligner-core/src/pwm/mod.rs:29-126, aligner-core/src/alignment.rs:45-79,
// bin/latent-repeat-search/engine/calc.rs:47-69, :107-137).
//
// The fill is the warp-per-item kernel of the pairwise aligner (kernels_warp.cuh) in its PWM view
// (common.cuh ChunkView::pwm_len): rows = window symbols, columns = profile positions, the 4 x L
// profile in shared memory.  Windows whose penalty speculation does not settle (or a profile with
// non-finite entries) go to the exact column-major kernel.  The traceback below differs from the
// pairwise one exactly where the reference does: no seed pair, position numbers instead of query
// symbols, arg-max (0,0) is an empty alignment.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "host_util.h"
#include "kernels_exact.cuh"
#include "kernels_trace.cuh"
#include "warp_args.h"

namespace alnr {

struct PwmTraceArgs {
  ChunkView cv;
  const uint32_t *D;
  DLayout dl;
  FillOut fo;
  uint64_t *coords;    // [pair*4]
  uint64_t *aln_off;   // [pair], absolute entry index (pair * aln_slot + ...)
  uint32_t *aln_len;   // [pair]
  uint32_t *numbered;  // chunk staging: [(pair - cv.first) * aln_slot + k]
  uint8_t *query;
  uint64_t aln_slot;
  double *freq;        // chunk staging: [(pair - cv.first) * 4 * L]
};

// One thread per window walks the 2-bit D scratch from the arg-max until Beginning
// (pwm/mod.rs:77-104) and writes the two strings backwards from the end of the window's slot
// (the reference pushes and reverses, :106-107).
static __global__ void __launch_bounds__(128) pwm_traceback_kernel(PwmTraceArgs a) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= a.cv.count) return;
  const uint32_t pair = a.cv.first + slot;
  const uint32_t L = a.cv.pwm_len;
  const uint64_t slot_end = ((uint64_t)slot + 1) * a.aln_slot;  // in the chunk staging buffers
  const uint64_t abs_base = (uint64_t)a.cv.first * a.aln_slot;  // staging entry 0 in the caller's arrays
  double *freq = a.freq ? a.freq + (size_t)slot * 4 * L : nullptr;
  if (freq) for (uint32_t i = 0; i < 4 * L; ++i) freq[i] = 0.0;
  if (a.fo.status[pair] != 0) {
    a.aln_len[pair] = 0;
    a.aln_off[pair] = abs_base + slot_end;
    for (int k = 0; k < 4; ++k) a.coords[4 * (size_t)pair + k] = 0;
    return;
  }
  const uint8_t *q, *t;
  uint32_t n, m;
  a.cv.pair_dims(pair, q, t, n, m);
  const uint32_t sy = a.fo.start[2 * (size_t)pair], sx = a.fo.start[2 * (size_t)pair + 1];
  uint32_t cy = sy, cx = sx;
  uint64_t pos = slot_end;
  const uint32_t rpw = a.dl.rows_per_word;
  const bool wr = a.numbered != nullptr && a.query != nullptr;
  auto emit = [&](uint32_t num, uint8_t sym) {
    --pos;
    if (wr) { a.numbered[pos] = num; a.query[pos] = sym; }
    if (freq && num != 0 && sym != kBlank) freq[(size_t)sym * L + (num - 1)] += 1.0;  // alignment.rs:58-62
  };
  for (;;) {
    if (cy == 0 || cx == 0) break;  // row 0 / column 0 stay Beginning (pwm/mod.rs:49-50)
    const uint32_t w = a.D[a.dl.word(slot, (cy - 1) / rpw, cx - 1)];
    const int d = a.dl.get(w, (cy - 1) % rpw, cx - 1);
    if (d == DIR_BEGINNING) break;
    if (d == DIR_TOP) { emit(0u, t[cy - 1]); --cy; }                 // :86-90
    else if (d == DIR_LEFT) { emit(cx, (uint8_t)kBlank); --cx; }      // :91-95
    else { emit(cx, t[cy - 1]); --cx; --cy; }                         // :96-101
  }
  a.aln_len[pair] = (uint32_t)(slot_end - pos);
  a.aln_off[pair] = abs_base + pos;
  uint64_t *c = a.coords + 4 * (size_t)pair;
  c[0] = cx + 1; c[1] = sx + 1; c[2] = cy + 1; c[3] = sy + 1;         // :118-121
}

// score-only calls: coords are not defined without the walk; report the arg-max as (x_to, y_to)
static __global__ void pwm_end_coords_kernel(FillOut fo, uint64_t *coords, uint32_t first, uint32_t count) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= count) return;
  const uint32_t pair = first + slot;
  uint64_t *c = coords + 4 * (size_t)pair;
  const bool ok = fo.status[pair] == 0;
  c[0] = 0; c[2] = 0;
  c[1] = ok ? fo.start[2 * (size_t)pair + 1] + 1 : 0;
  c[3] = ok ? fo.start[2 * (size_t)pair] + 1 : 0;
}

}  // namespace alnr

using namespace alnr;

struct alnr_pwm_seq {
  int device = 0;
  uint64_t len = 0;
  DevBuf<uint8_t> d_seq;
  DevBuf<int> d_flag;  // [0] != 0: some symbol >= 4 (windows are then checked one by one)
};

static thread_local int t_pwm_numeric = 0;

extern "C" {

int alnr_pwm_seq_create(const uint8_t *seq, uint64_t len, int device, alnr_pwm_seq **h) {
  if (!h) return ALNR_INVALID_ARGUMENT;
  *h = nullptr;
  if (len && !seq) return ALNR_INVALID_ARGUMENT;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    (void)cudaGetLastError();
    set_error_detail("no usable CUDA device (this library has no CPU fallback)");
    return ALNR_CUDA_ERROR;
  }
  if (device < 0) device = 0;
  if (device >= ndev) return ALNR_INVALID_ARGUMENT;
  ALNR_CUDA(cudaSetDevice(device));
  alnr_pwm_seq *s = new (std::nothrow) alnr_pwm_seq();
  if (!s) return ALNR_OUT_OF_MEMORY;
  s->device = device;
  s->len = len;
  auto fail = [&](int st) { delete s; return st; };
  cudaError_t e = s->d_seq.alloc(len + 16);
  if (e == cudaSuccess) e = s->d_flag.alloc(4);
  if (e == cudaSuccess) e = cudaMemset(s->d_flag.p, 0, 4 * sizeof(int));
  if (e == cudaSuccess && len) e = cudaMemcpy(s->d_seq.p, seq, len, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return fail(cuda_fail(e, "upload sequence", __FILE__, __LINE__));
  if (len) {
    any_symbol_out_of_range_kernel<<<148 * 8, 256>>>(s->d_seq.p, len, 4, s->d_flag.p);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return fail(cuda_fail(e, "symbol sweep", __FILE__, __LINE__));
  }
  *h = s;
  return ALNR_OK;
}

uint64_t alnr_pwm_seq_len(const alnr_pwm_seq *h) { return h ? h->len : 0; }

void alnr_pwm_seq_destroy(alnr_pwm_seq *h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  delete h;
}

int alnr_pwm_last_numeric(void) { return t_pwm_numeric; }

}  // extern "C"

namespace {

// Integer domain: every profile entry and both penalties are integers and no |H| can reach 2^30.
int pwm_classify(const alnr_pwm_params *p, uint32_t m_max) {
  if (p->numeric == ALNR_NUM_F64 || p->numeric == ALNR_NUM_I32) return p->numeric;
  auto integral = [](double v) { return std::fabs(v) <= 1048576.0 && v == std::floor(v); };
  double amax = std::max(std::fabs(p->del), std::fabs(p->ext));
  bool ok = integral(p->del) && integral(p->ext);
  for (size_t i = 0; ok && i < 4 * (size_t)p->length; ++i) {
    ok = integral(p->pwm[i]);
    amax = std::max(amax, std::fabs(p->pwm[i]));
  }
  if (ok && amax * ((double)p->length + (double)m_max + 2.0) < 1073741824.0) return ALNR_NUM_I32;
  return ALNR_NUM_F64;
}

template <typename T> int pwm_warp_launch(cudaStream_t s, const WarpArgs<T> &a, bool mats);
template <> int pwm_warp_launch<int>(cudaStream_t s, const WarpArgs<int> &a, bool mats) { return warp_launch_i32(s, a, true, mats); }
template <> int pwm_warp_launch<double>(cudaStream_t s, const WarpArgs<double> &a, bool mats) { return warp_launch_f64(s, a, true, mats); }

template <typename T>
int pwm_run(const alnr_pwm_params *p, alnr_pwm_seq *seq, const alnr_windows *w, alnr_pwm_out *out,
            uint32_t m_max) {
  const uint32_t L = p->length;
  const uint64_t count = w->count;
  const bool want_aln = out->numbered || out->query || out->freq;
  const bool mats = out->H || out->D;
  const bool stateful = p->del != p->ext;
  bool finite = std::fabs(p->del) < 1e290 && std::fabs(p->ext) < 1e290;
  for (size_t i = 0; i < 4 * (size_t)L; ++i) finite = finite && std::fabs(p->pwm[i]) < 1e290;
  const bool use_warp = sizeof(T) == 4 || finite;  // FP64 warp kernel: compare-select max needs finite inputs
  const uint32_t mm = std::max(m_max, 1u);
  const DLayout dl = use_warp ? DLayout::make_skewed(mats ? 8 : warp_pick_rows(mm), mm, L)
                              : DLayout::make(16, mm, L, false);
  const uint32_t nzw = (L + 31) / 32;
  const uint64_t row_words = (uint64_t)std::max(m_max, L) + 1;
  const uint64_t slot = out->aln_slot;
  const uint64_t per_item = (want_aln || mats ? dl.words_for(32) * 4 / 32 : 0) + sizeof(T) * row_words +
                            (stateful ? 8ull * nzw + 4 : 4) + 64 +
                            ((out->numbered || out->query) ? 5 * slot : 0) + (out->freq ? 32ull * L : 0);
  size_t free_b = 0, total_b = 0;
  ALNR_CUDA(cudaMemGetInfo(&free_b, &total_b));
  const uint64_t budget = std::max<uint64_t>(std::min<uint64_t>(free_b / 3, 24ull << 30), 64ull << 20);
  uint64_t chunk = std::max<uint64_t>(budget / per_item / 32 * 32, 32);
  chunk = std::min<uint64_t>(chunk, std::min<uint64_t>((count + 31) / 32 * 32, 1u << 22));

  cudaStream_t stream;
  ALNR_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamSynchronize(s); cudaStreamDestroy(s); } } guard{stream};

  DevBuf<T> d_pwm, d_hcol;
  DevBuf<uint64_t> d_starts, d_coords, d_alnoff;
  DevBuf<uint32_t> d_lens, d_start, d_alnlen, d_D, d_wz, d_num;
  DevBuf<uint8_t> d_qry, d_Dout;
  DevBuf<int32_t> d_status;
  DevBuf<double> d_f, d_corner, d_freq, d_Hout;
  // declared after the buffers, so destroyed first: nothing returns to the pool while kernels of an
  // early return are still queued
  struct SyncGuard { cudaStream_t s; ~SyncGuard() { cudaStreamSynchronize(s); (void)cudaGetLastError(); } } sync_guard{stream};
  std::vector<T> hp(4 * (size_t)L);
  for (size_t i = 0; i < hp.size(); ++i) hp[i] = (T)p->pwm[i];
  ALNR_CUDA(d_pwm.alloc(hp.size()));
  ALNR_CUDA(cudaMemcpyAsync(d_pwm.p, hp.data(), hp.size() * sizeof(T), cudaMemcpyHostToDevice, stream));
  if (w->starts) {
    ALNR_CUDA(d_starts.alloc(count)); ALNR_CUDA(d_lens.alloc(count));
    ALNR_CUDA(cudaMemcpyAsync(d_starts.p, w->starts, count * 8, cudaMemcpyHostToDevice, stream));
    ALNR_CUDA(cudaMemcpyAsync(d_lens.p, w->lens, count * 4, cudaMemcpyHostToDevice, stream));
  }
  ALNR_CUDA(d_f.alloc(count)); ALNR_CUDA(d_corner.alloc(count)); ALNR_CUDA(d_status.alloc(count));
  ALNR_CUDA(d_start.alloc(2 * count)); ALNR_CUDA(d_coords.alloc(4 * count));
  ALNR_CUDA(d_alnoff.alloc(count)); ALNR_CUDA(d_alnlen.alloc(count));
  ALNR_CUDA(d_hcol.alloc((chunk + 31) / 32 * 32 * row_words));
  ALNR_CUDA(d_wz.alloc(warp_scratch_words(chunk, nzw, stateful)));
  if (want_aln || mats) ALNR_CUDA(d_D.alloc(dl.words_for(chunk)));
  if (out->numbered || out->query) { ALNR_CUDA(d_num.alloc(chunk * slot)); ALNR_CUDA(d_qry.alloc(chunk * slot)); }
  if (out->freq) ALNR_CUDA(d_freq.alloc(chunk * 4 * L));
  if (mats) {
    ALNR_CUDA(d_Hout.alloc(((size_t)m_max + 1) * ((size_t)L + 1)));
    ALNR_CUDA(d_Dout.alloc(((size_t)m_max + 1) * ((size_t)L + 1)));
  }
  const size_t smem_exact = 4 * (size_t)L * sizeof(T);
  if (smem_exact > 48 * 1024)
    ALNR_CUDA(cudaFuncSetAttribute(fill_exact_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_exact));

  auto d2h = [&](void *dst, const void *src, size_t bytes) -> cudaError_t {
    if (!dst || !bytes) return cudaSuccess;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream);
  };
  for (uint64_t base = 0; base < count; base += chunk) {
    const uint32_t cnt = (uint32_t)std::min<uint64_t>(chunk, count - base);
    ChunkView cv{seq->d_seq.p, seq->d_seq.p, nullptr, nullptr, nullptr, (uint32_t)base, cnt};
    cv.pwm_len = L;
    cv.win_starts = w->starts ? d_starts.p : nullptr;
    cv.win_lens = w->starts ? d_lens.p : nullptr;
    cv.win_first = w->first; cv.win_step = w->step; cv.win_size = w->window; cv.seq_len = seq->len;
    FillOut fo{d_f.p, d_corner.p, d_start.p, d_status.p};
    validate_kernel<<<(cnt + 127) / 128, 128, 0, stream>>>(cv, 4, d_status.p, seq->d_flag.p); ALNR_LAUNCHED();
    uint32_t *D = (want_aln || mats) ? d_D.p : nullptr;
    ExactArgs<T> e;
    e.cv = cv; e.mode = ALNR_LOCAL; e.V = 4; e.del = (T)p->del; e.ext = (T)p->ext;
    e.matrix = d_pwm.p; e.hcol = d_hcol.p; e.hrows = m_max + 1; e.D = D; e.dl = dl; e.fo = fo;
    e.H_out = mats ? d_Hout.p : nullptr; e.D_out = mats ? d_Dout.p : nullptr;
    if (use_warp) {
      WarpArgs<T> a;
      a.cv = cv; a.V = 4; a.del = (T)p->del; a.ext = (T)p->ext; a.stateful = stateful;
      a.matrix = d_pwm.p; a.item_matrices = nullptr; a.D = D; a.dl = dl; a.fo = fo;
      a.rowbuf = d_hcol.p; a.rowbuf_stride = L + 1;
      a.nzw = nzw;
      warp_scratch_bind(a, d_wz.p, chunk, !mats);
      a.H_out = mats ? d_Hout.p : nullptr; a.D_out = mats ? d_Dout.p : nullptr;
      ALNR_CUDA(cudaMemsetAsync(d_wz.p, 0, 8 * sizeof(uint32_t), stream));
      int st = pwm_warp_launch<T>(stream, a, mats);
      if (st) return st;
      if (stateful) {  // windows whose speculation did not settle
        e.list = a.fail_list; e.list_count = a.fail_count;
        fill_exact_kernel<T><<<(cnt + 127) / 128, 128, smem_exact, stream>>>(e); ALNR_LAUNCHED();
      }
    } else {
      fill_exact_kernel<T><<<(cnt + 127) / 128, 128, smem_exact, stream>>>(e); ALNR_LAUNCHED();
    }
    if (want_aln) {
      PwmTraceArgs ta{cv, D, dl, fo, d_coords.p, d_alnoff.p, d_alnlen.p,
                      (out->numbered || out->query) ? d_num.p : nullptr, (out->numbered || out->query) ? d_qry.p : nullptr,
                      slot, out->freq ? d_freq.p : nullptr};
      pwm_traceback_kernel<<<(cnt + 127) / 128, 128, 0, stream>>>(ta); ALNR_LAUNCHED();
    } else {
      pwm_end_coords_kernel<<<(cnt + 127) / 128, 128, 0, stream>>>(fo, d_coords.p, (uint32_t)base, cnt); ALNR_LAUNCHED();
    }
    ALNR_CUDA(d2h(out->numbered ? out->numbered + base * slot : nullptr, d_num.p, (size_t)cnt * slot * 4));
    ALNR_CUDA(d2h(out->query ? out->query + base * slot : nullptr, d_qry.p, (size_t)cnt * slot));
    ALNR_CUDA(d2h(out->freq ? out->freq + base * 4 * L : nullptr, d_freq.p, (size_t)cnt * 4 * L * 8));
    ALNR_CUDA(cudaStreamSynchronize(stream));  // the staging buffers are reused by the next chunk
    static const bool timing = getenv("ALNR_TIMING") != nullptr;
    if (timing && use_warp && stateful) {  // exact-kernel list, mispredicted, left after the verify kernel
      uint32_t hc[4] = {0, 0, 0, 0};
      ALNR_CUDA(cudaMemcpy(hc, d_wz.p, sizeof(hc), cudaMemcpyDeviceToHost));
      std::fprintf(stderr, "[alnr] pwm speculation: %u windows, mispredicted %u, second launch %u, exact kernel %u\n",
                   cnt, hc[1], hc[2], hc[0]);
    }
  }
  ALNR_CUDA(d2h(out->f, d_f.p, count * 8));
  ALNR_CUDA(d2h(out->coords, d_coords.p, count * 32));
  ALNR_CUDA(d2h(out->status, d_status.p, count * 4));
  if (want_aln) {
    ALNR_CUDA(d2h(out->aln_off, d_alnoff.p, count * 8));
    ALNR_CUDA(d2h(out->aln_len, d_alnlen.p, count * 4));
  }
  if (mats) {
    const size_t cells = ((size_t)m_max + 1) * ((size_t)L + 1);
    ALNR_CUDA(d2h(out->H, d_Hout.p, cells * 8));
    ALNR_CUDA(d2h(out->D, d_Dout.p, cells));
  }
  ALNR_CUDA(cudaStreamSynchronize(stream));
  return ALNR_OK;
}

}  // namespace

extern "C" int alnr_pwm_align_windows(const alnr_pwm_params *p, alnr_pwm_seq *seq, const alnr_windows *w,
                                      alnr_pwm_out *out) {
  if (!p || !seq || !w || !out) return ALNR_INVALID_ARGUMENT;
  if (p->rows != 4) return ALNR_MATRIX_SHAPE_ERROR;  // pwm/mod.rs:40-42
  if (p->length && !p->pwm) return ALNR_INVALID_ARGUMENT;
  if (p->device >= 0 && p->device != seq->device) return ALNR_INVALID_ARGUMENT;
  if (!(std::fabs(p->del) < 1e290 && std::fabs(p->ext) < 1e290)) return ALNR_INVALID_ARGUMENT;  // NaN / inf penalties
  const uint64_t count = w->count;
  if (count == 0) return ALNR_OK;
  if (count >= (1ull << 31)) return ALNR_INVALID_ARGUMENT;
  if ((w->starts == nullptr) != (w->lens == nullptr)) return ALNR_INVALID_ARGUMENT;
  if ((out->H || out->D) && count != 1) return ALNR_INVALID_ARGUMENT;
  const uint32_t L = p->length;
  if (4ull * L * 8 > 200 * 1024) {
    set_error_detail("PWM longer than 6400 columns does not fit shared memory");
    return ALNR_INVALID_ARGUMENT;
  }
  uint32_t m_max = 0;
  if (w->starts) {
    for (uint64_t i = 0; i < count; ++i) {
      if (w->starts[i] > seq->len || w->lens[i] > seq->len - w->starts[i]) return ALNR_INVALID_ARGUMENT;
      m_max = std::max(m_max, w->lens[i]);
    }
  } else if (w->first < seq->len) {
    m_max = (uint32_t)std::min<uint64_t>(w->window, seq->len - w->first);
  }
  if (m_max >= (1u << 30)) return ALNR_INVALID_ARGUMENT;
  const bool want_strings = out->numbered || out->query;
  if (want_strings && (!out->numbered || !out->query || !out->aln_off || !out->aln_len)) return ALNR_INVALID_ARGUMENT;
  if (want_strings && out->aln_slot < (uint64_t)L + m_max) return ALNR_INVALID_ARGUMENT;
  ALNR_CUDA(cudaSetDevice(seq->device));
  if (L == 0) {  // dim = (m + 1, 1): nothing to fill, arg-max (0,0) (pwm/mod.rs:46-50, :76)
    for (uint64_t i = 0; i < count; ++i) {
      if (out->f) out->f[i] = 0.0;
      if (out->status) out->status[i] = 0;
      if (out->coords) for (int k = 0; k < 4; ++k) out->coords[4 * i + k] = 1;
      if (out->aln_len) out->aln_len[i] = 0;
      if (out->aln_off) out->aln_off[i] = (i + 1) * out->aln_slot;
    }
    if (out->H) for (uint64_t y = 0; y <= m_max; ++y) out->H[y] = 0.0;
    if (out->D) for (uint64_t y = 0; y <= m_max; ++y) out->D[y] = ALNR_DIR_BEGINNING;
    return ALNR_OK;
  }
  const int numeric = pwm_classify(p, m_max);
  if (numeric == ALNR_NUM_I32) {  // a forced integer domain must be exact
    bool ok = p->del == std::floor(p->del) && p->ext == std::floor(p->ext);
    for (size_t i = 0; ok && i < 4 * (size_t)L; ++i) ok = p->pwm[i] == std::floor(p->pwm[i]) && std::fabs(p->pwm[i]) < 1e9;
    if (!ok) return ALNR_INVALID_ARGUMENT;
  }
  t_pwm_numeric = numeric;
  return numeric == ALNR_NUM_I32 ? pwm_run<int>(p, seq, w, out, m_max) : pwm_run<double>(p, seq, w, out, m_max);
}
