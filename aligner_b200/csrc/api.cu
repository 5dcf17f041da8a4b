// api.cu — the extern "C" boundary declared in include/aligner_b200.h.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "batch.h"

namespace alnr { const char *get_error_detail(); }
using alnr::alnr_batch_impl;

struct alnr_batch_handle {
  alnr_batch_impl impl;
};

extern "C" {

const char *alnr_strerror(int status) {
  switch (status) {
    case ALNR_OK: return "ok";
    case ALNR_PROTEIN_NOT_FOUND: return "ProteinNotFound";
    case ALNR_CHAR_IS_NOT_MATCHABLE: return "CharIsNotMatchable";
    case ALNR_UNNECESSARY_ARGUMENT: return "UnnecessaryArgument";
    case ALNR_MISSING_ARGUMENT: return "MissingArgument";
    case ALNR_RESULT_IS_EMPTY: return "ResultIsEmpty";
    case ALNR_CALCULATION_ERROR: return "CalculationError";
    case ALNR_VALIDATION_ERROR: return "ValidationError";
    case ALNR_MATRIX_SHAPE_ERROR: return "MatrixShapeError";
    case ALNR_WRONG_MATRIX: return "WrongMatrixSpecified";
    case ALNR_EMPTY_INPUT: return "empty sequence (reference panics: unwrap on None)";
    case ALNR_NO_POSITIVE_CELL: return "no positive cell in local matrix (reference panics: usize underflow)";
    case ALNR_SYMBOL_OUT_OF_RANGE: return "symbol code outside the alphabet (reference panics: index out of bounds)";
    case ALNR_CUDA_ERROR: return "CUDA error";
    case ALNR_INVALID_ARGUMENT: return "invalid argument";
    case ALNR_OUT_OF_MEMORY: return "out of device memory";
    default: return "unknown status";
  }
}

const char *alnr_last_error_detail(void) { return alnr::get_error_detail(); }

int alnr_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
  return n;
}

const char *alnr_build_info(void) { return "aligner_b200 sm_100a cuda " __DATE__; }

uint64_t alnr_kernel_launch_count(void) { return alnr::g_launches.load(); }

int alnr_batch_create(const alnr_params *p, const alnr_batch *b, alnr_batch_handle **h) {
  if (!h) return ALNR_INVALID_ARGUMENT;
  *h = nullptr;
  alnr_batch_handle *hh = new (std::nothrow) alnr_batch_handle();
  if (!hh) return ALNR_OUT_OF_MEMORY;
  int st = hh->impl.init(p, b, false);
  if (st) { delete hh; return st; }
  *h = hh;
  return ALNR_OK;
}

int alnr_batch_create_compact(const alnr_params *p, const alnr_batch *b, alnr_batch_handle **h) {
  if (!h) return ALNR_INVALID_ARGUMENT;
  *h = nullptr;
  alnr_batch_handle *hh = new (std::nothrow) alnr_batch_handle();
  if (!hh) return ALNR_OUT_OF_MEMORY;
  hh->impl.moves_out = true;
  int st = hh->impl.init(p, b, false);
  if (st) { delete hh; return st; }
  *h = hh;
  return ALNR_OK;
}

uint64_t alnr_moves_capacity(const alnr_batch *b) {
  if (!b || !b->q_off || !b->t_off) return 0;
  uint64_t w = 0;
  for (uint64_t i = 0; i < b->count; ++i)
    w += (b->q_off[i + 1] - b->q_off[i] + b->t_off[i + 1] - b->t_off[i] + 15) >> 4;
  return w + 4;
}

int alnr_batch_create_scores(const alnr_params *p, const alnr_batch *b, alnr_batch_handle **h) {
  if (!h) return ALNR_INVALID_ARGUMENT;
  *h = nullptr;
  alnr_batch_handle *hh = new (std::nothrow) alnr_batch_handle();
  if (!hh) return ALNR_OUT_OF_MEMORY;
  hh->impl.score_only = true;
  int st = hh->impl.init(p, b, false);
  if (st) { delete hh; return st; }
  *h = hh;
  return ALNR_OK;
}

int alnr_batch_download_scores(alnr_batch_handle *h, double *f, double *corner, uint32_t *end_yx,
                               int32_t *status) {
  return h ? h->impl.download_scores(f, corner, end_yx, status) : ALNR_INVALID_ARGUMENT;
}

int alnr_score_batch(const alnr_params *p, const alnr_batch *b, double *f, double *corner,
                     uint32_t *end_yx, int32_t *status) {
  alnr_batch_handle *hh = nullptr;
  int st = alnr_batch_create_scores(p, b, &hh);
  if (st) return st;
  st = hh->impl.run();
  if (!st) st = hh->impl.download_scores(f, corner, end_yx, status);
  alnr_batch_destroy(hh);
  return st;
}

int alnr_batch_run(alnr_batch_handle *h) { return h ? h->impl.run() : ALNR_INVALID_ARGUMENT; }

int alnr_batch_sync(alnr_batch_handle *h) {
  if (!h) return ALNR_INVALID_ARGUMENT;
  ALNR_CUDA(cudaSetDevice(h->impl.device));
  ALNR_CUDA(cudaStreamSynchronize(h->impl.stream));
  return ALNR_OK;
}

int alnr_batch_download(alnr_batch_handle *h, alnr_batch_out *out) {
  return h ? h->impl.download(out) : ALNR_INVALID_ARGUMENT;
}

void *alnr_batch_stream(alnr_batch_handle *h) { return h ? (void *)h->impl.stream : nullptr; }

int alnr_batch_run_timed(alnr_batch_handle *h, int steps, float *ms_total, float *fill_ms_total) {
  if (!h || steps < 1) return ALNR_INVALID_ARGUMENT;
  alnr_batch_impl &B = h->impl;
  ALNR_CUDA(cudaSetDevice(B.device));
  ALNR_CUDA(cudaStreamSynchronize(B.stream));
  B.profile = fill_ms_total != nullptr;
  B.fill_events_used = 0;
  ALNR_CUDA(cudaEventRecord(B.ev0, B.stream));
  for (int s = 0; s < steps; ++s) {
    int st = B.run();
    if (st) { B.profile = false; return st; }
  }
  ALNR_CUDA(cudaEventRecord(B.ev1, B.stream));
  ALNR_CUDA(cudaEventSynchronize(B.ev1));
  float t = 0.f;
  ALNR_CUDA(cudaEventElapsedTime(&t, B.ev0, B.ev1));
  if (ms_total) *ms_total = t;
  if (B.profile) {
    int st = B.collect_fill_ms();
    B.profile = false;
    if (st) return st;
    *fill_ms_total = B.last_fill_ms;
  }
  return ALNR_OK;
}

int alnr_batch_numeric(alnr_batch_handle *h) { return h ? h->impl.numeric : 0; }
uint64_t alnr_batch_rerun_count(alnr_batch_handle *h) { return h ? h->impl.rerun_total : 0; }
void alnr_batch_destroy(alnr_batch_handle *h) {
  if (!h) return;
  cudaSetDevice(h->impl.device);
  if (h->impl.stream) cudaStreamSynchronize(h->impl.stream);  // buffers go back to the pool
  delete h;
}

int alnr_align_batch(const alnr_params *p, const alnr_batch *b, alnr_batch_out *out) {
  if (!out) return ALNR_INVALID_ARGUMENT;
  if (out->moves && (out->q_aln || out->t_aln || out->aln_off)) return ALNR_INVALID_ARGUMENT;
  alnr_batch_handle *hh = new (std::nothrow) alnr_batch_handle();
  if (!hh) return ALNR_OUT_OF_MEMORY;
  alnr_batch_impl &B = hh->impl;
  B.moves_out = out->moves != nullptr;
  B.want_coords = out->coords != nullptr;
  // uniform batches stream chunk by chunk (copies overlap compute); init() finds out which
  B.resident = false;
  B.ramp = true;
  B.max_chunk = 620000;  // ~4 waves of the strip kernel: short pipeline fill / drain
  if (const char *mc = std::getenv("ALNR_STREAM_CHUNK")) B.max_chunk = (uint32_t)std::max(1L, std::atol(mc));
  static const bool timing = std::getenv("ALNR_TIMING") != nullptr;
  const auto t0 = std::chrono::steady_clock::now();
  int st = B.init(p, b, false);
  const auto t1 = std::chrono::steady_clock::now();
  if (!st) {
    if (B.uniform && (!B.moves_out || B.moves_direct)) {
      st = B.stream_all(b, out);
      if (timing)
        std::fprintf(stderr, "[alnr] align_batch: init %.1f ms, stream_all %.1f ms, %zu chunks\n",
                     std::chrono::duration<double, std::milli>(t1 - t0).count(),
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count(),
                     B.chunks.size());
    } else {
      cudaError_t e = cudaSuccess;
      if (B.count && B.packed_in) {
        e = cudaMemcpyAsync(B.d_q2raw.p, b->q2, B.q2_words * 4, cudaMemcpyHostToDevice, B.stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(B.d_t2raw.p, b->t2, B.t2_words * 4, cudaMemcpyHostToDevice, B.stream);
      } else if (B.count) {
        e = cudaMemcpyAsync(B.d_q.p, b->q, B.q_total, cudaMemcpyHostToDevice, B.stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(B.d_t.p, b->t, B.t_total, cudaMemcpyHostToDevice, B.stream);
      }
      if (e != cudaSuccess) st = alnr::cuda_fail(e, "H2D", __FILE__, __LINE__);
      if (!st) st = B.run();
      if (!st) st = B.download(out);
    }
  }
  alnr_batch_destroy(hh);
  return st;
}

}  // extern "C"

namespace {
// Device -> pageable host memory through two pinned 4 MB buffers per calling thread: the copy engine
// fills one while the host empties the other (a plain cudaMemcpy into malloc'd memory reached
// ~1 GB/s for the 10 MB H matrix of a 1231 x 1020 pair).
struct PinnedPair {
  void *buf[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  static constexpr size_t kBytes = 4u << 20;
  bool ready() {
    if (buf[0]) return true;
    for (int k = 0; k < 2; ++k)
      if (cudaMallocHost(&buf[k], kBytes) != cudaSuccess ||
          cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming) != cudaSuccess) {
        (void)cudaGetLastError();
        release();
        return false;
      }
    return true;
  }
  void release() {
    for (int k = 0; k < 2; ++k) {
      if (buf[k]) cudaFreeHost(buf[k]);
      if (ev[k]) cudaEventDestroy(ev[k]);
      buf[k] = nullptr; ev[k] = nullptr;
    }
    (void)cudaGetLastError();
  }
  ~PinnedPair() { release(); }
};

cudaError_t d2h_staged(void *dst, const void *src, size_t bytes, cudaStream_t stream) {
  static thread_local PinnedPair pp;
  if (bytes < (256u << 10) || !pp.ready()) return cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost);
  const size_t nchunks = (bytes + PinnedPair::kBytes - 1) / PinnedPair::kBytes;
  auto len = [&](size_t c) { return std::min(PinnedPair::kBytes, bytes - c * PinnedPair::kBytes); };
  cudaError_t e = cudaSuccess;
  for (size_t c = 0; c <= nchunks && e == cudaSuccess; ++c) {
    if (c < nchunks) {
      e = cudaMemcpyAsync(pp.buf[c & 1], (const char *)src + c * PinnedPair::kBytes, len(c), cudaMemcpyDeviceToHost, stream);
      if (e == cudaSuccess) e = cudaEventRecord(pp.ev[c & 1], stream);
    }
    if (c > 0 && e == cudaSuccess) {
      e = cudaEventSynchronize(pp.ev[(c - 1) & 1]);
      if (e == cudaSuccess) std::memcpy((char *)dst + (c - 1) * PinnedPair::kBytes, pp.buf[(c - 1) & 1], len(c - 1));
    }
  }
  return e;
}
}  // namespace

extern "C" {

int alnr_align_pair(const alnr_params *p, const uint8_t *q, uint64_t n, const uint8_t *t,
                    uint64_t m, int want_matrices, alnr_pair_result *out) {
  if (!out || !p) return ALNR_INVALID_ARGUMENT;
  std::memset(out, 0, sizeof(*out));
  if (n == 0 || m == 0) return ALNR_EMPTY_INPUT;
  if (!q || !t) return ALNR_INVALID_ARGUMENT;
  uint64_t qo[2] = {0, n}, to[2] = {0, m};
  alnr_batch b{q, qo, t, to, 1};
  alnr_batch_handle hh;
  int st = hh.impl.init(p, &b, want_matrices != 0);
  if (st) return st;
  st = hh.impl.run();
  if (st) return st;
  const size_t cap = n + m + 1, cells = (n + 1) * (m + 1);
  uint8_t *qa = (uint8_t *)std::malloc(cap), *ta = (uint8_t *)std::malloc(cap);
  double *H = want_matrices ? (double *)std::malloc(cells * sizeof(double)) : nullptr;
  uint8_t *D = want_matrices ? (uint8_t *)std::malloc(cells) : nullptr;
  auto fail = [&](int code) { std::free(qa); std::free(ta); std::free(H); std::free(D); return code; };
  if (!qa || !ta || (want_matrices && (!H || !D))) return fail(ALNR_OUT_OF_MEMORY);
  uint64_t off = 0;
  uint32_t len = 0;
  int32_t pst = 0;
  alnr_batch_out bo{&out->f, &out->corner_score, out->coords, &off, &len, qa, ta, &pst};
  st = hh.impl.download(&bo);
  if (st) return fail(st);
  if (pst) return fail(pst);
  if (want_matrices) {
    cudaError_t e = d2h_staged(H, hh.impl.d_Hout.p, cells * sizeof(double), hh.impl.stream);
    if (e == cudaSuccess) e = d2h_staged(D, hh.impl.d_Dout.p, cells, hh.impl.stream);
    if (e != cudaSuccess) return fail(alnr::cuda_fail(e, "copy H/D", __FILE__, __LINE__));
  }
  std::memmove(qa, qa + off, len);  // slot is right-aligned (aligner_b200.h)
  std::memmove(ta, ta + off, len);
  out->q_aln = qa; out->t_aln = ta; out->aln_len = len; out->H = H; out->D = D;
  out->numeric = hh.impl.numeric;
  return ALNR_OK;
}

void alnr_free_pair_result(alnr_pair_result *r) {
  if (!r) return;
  std::free(r->q_aln); std::free(r->t_aln); std::free(r->H); std::free(r->D);
  r->q_aln = r->t_aln = nullptr; r->H = nullptr; r->D = nullptr;
}

void *alnr_host_alloc(uint64_t bytes) {
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
  return p;
}
void alnr_host_free(void *p) { if (p) cudaFreeHost(p); }

}  // extern "C"
