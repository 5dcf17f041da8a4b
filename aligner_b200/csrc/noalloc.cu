// noalloc.cu — alnr_score_batch_noalloc: a score-only batch call that allocates NOTHING.
//
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  It is a variant of fill_exact_kernel (kernels_exact.cuh: one pair per thread, cells in the
// reference's own x-outer / y-inner order, simple/mod.rs:74-97, :187-210), score-only, whose scratch
// — the previous score column of every pair — lives in a file-scope static __device__ array instead
// of a pool buffer.  Inputs and outputs are DEVICE pointers supplied by the caller; the substitution
// matrix travels by value in the kernel arguments, so the call issues no copy, no memset, no
// cudaMalloc / cudaFree / cudaMallocHost and no allocating library routine: one kernel launch.
//
// The array costs ALNR_NOALLOC_SCRATCH_BYTES (16 MiB) on every GPU that loads the library, until
// exit.  EVERY call on a device shares it and it keeps the last call's contents: nothing in it has
// to start at zero (each thread writes column 0 of its own slice before reading it), and two calls
// that overlap on the same GPU (different streams or host threads) RACE — serialise them.
#include <cmath>
#include <cstdio>
#include <mutex>

#include "host_util.h"
#include "kernels_exact.cuh"

namespace alnr {

constexpr size_t kNoAllocBytes = ALNR_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

struct NoAllocArgs {
  const uint8_t *q, *t;          // device: concatenated symbol codes
  const uint64_t *q_off, *t_off; // device: count + 1 offsets
  uint32_t count;
  uint32_t n_cap, m_cap;         // the lengths the caller promised (scratch is sized for m_cap)
  int mode, V;
  double del, ext;
  unsigned char *scratch;        // address of g_noalloc_scratch on this device (cudaGetSymbolAddress)
  double *f, *corner;            // device outputs, any may be nullptr
  uint32_t *end_yx;
  int32_t *status;
  double matrix[32 * 32];        // V * V, [t][q], by value
};

// grid: ceil(count / 128) x 128 threads; dynamic smem: V * V * sizeof(T)
template <typename T>
__global__ void __launch_bounds__(128) score_exact_static_kernel(const __grid_constant__ NoAllocArgs a) {
  extern __shared__ unsigned char smem_raw[];
  T *smat = reinterpret_cast<T *>(smem_raw);
  for (int i = threadIdx.x; i < a.V * a.V; i += blockDim.x) smat[i] = (T)a.matrix[i];
  __syncthreads();
  const uint32_t pair = blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= a.count) return;
  const uint64_t q0 = a.q_off[pair], t0 = a.t_off[pair];
  const uint64_t nn = a.q_off[pair + 1] - q0, mm = a.t_off[pair + 1] - t0;
  auto finish = [&](int st, double f, double corner, uint32_t y, uint32_t x) {
    if (a.status) a.status[pair] = st;
    if (a.f) a.f[pair] = f;
    if (a.corner) a.corner[pair] = corner;
    if (a.end_yx) { a.end_yx[2 * (size_t)pair] = y; a.end_yx[2 * (size_t)pair + 1] = x; }
  };
  // what the pool-backed path finds out in validate_kernel, plus the promise about the lengths: a
  // longer target would run past this thread's slice of the shared scratch
  if (nn == 0 || mm == 0) return finish(ALNR_EMPTY_INPUT, 0.0, 0.0, 0, 0);
  if (nn > a.n_cap || mm > a.m_cap) return finish(ALNR_INVALID_ARGUMENT, 0.0, 0.0, 0, 0);
  const uint32_t n = (uint32_t)nn, m = (uint32_t)mm;
  const uint8_t *q = a.q + q0, *t = a.t + t0;
  for (uint32_t i = 0; i < n; ++i) if (q[i] >= a.V) return finish(ALNR_SYMBOL_OUT_OF_RANGE, 0.0, 0.0, 0, 0);
  for (uint32_t i = 0; i < m; ++i) if (t[i] >= a.V) return finish(ALNR_SYMBOL_OUT_OF_RANGE, 0.0, 0.0, 0, 0);

  const bool glob = a.mode == ALNR_GLOBAL;
  const T del = (T)a.del, ext = (T)a.ext;
  const uint32_t group = pair >> 5, lane = pair & 31;
  // hc[y * 32]: previous column of this pair; a warp's 32 pairs interleave -> coalesced
  T *hc = reinterpret_cast<T *>(a.scratch) + (uint64_t)group * (a.m_cap + 1) * 32 + lane;
  for (uint32_t y = 0; y <= m; ++y) {  // column 0 (simple/mod.rs:64-67, :70; zeros in local mode)
    T v = T(0);
    if (glob && y > 0) v = (y == m) ? -(T(m) + T(1)) * del : -T(y) * del;
    hc[(size_t)y * 32] = v;
  }
  T pen = del, best = T(0), last = T(0);
  uint32_t by = 0, bx = 0;
  for (uint32_t x = 1; x <= n; ++x) {
    const int qs = q[x - 1];
    T hup = T(0);
    if (glob) hup = (x == n) ? -(T(n) + T(1)) * del : -T(x) * del;  // simple/mod.rs:59-62, :69
    T diag = hc[0];
    hc[0] = hup;
    for (uint32_t y = 1; y <= m; ++y) {
      const T left = hc[(size_t)y * 32];
      const T top = nsub(hup, pen), lft = nsub(left, pen), dg = nadd(diag, smat[(int)t[y - 1] * a.V + qs]);
      const T mx = Num<T>::max3(top, lft, dg);                    // enums.rs:19
      const bool beginning = !glob && Num<T>::is_zero(mx);        // enums.rs:37
      pen = beginning ? del : ext;                                // simple/mod.rs:88-92, :201-205
      diag = left;
      hup = mx;
      hc[(size_t)y * 32] = mx;
      // ndarray-stats argmax: first maximum in row-major order (simple/mod.rs:212)
      if (!glob && (mx > best || (mx == best && y < by))) { best = mx; by = y; bx = x; }
    }
    last = hup;
  }
  if (glob) return finish(ALNR_OK, 0.0, (double)last, m, n);      // simple/mod.rs:139
  finish((by == 0 || bx == 0) ? ALNR_NO_POSITIVE_CELL : ALNR_OK, (double)best, (double)last, by, bx);
}

// Device address of the static array, looked up once per device with that device current.  In host
// code `g_noalloc_scratch` is only a stand-in for the symbol, never a device address.
static int noalloc_scratch_ptr(int device, unsigned char **out) {
  static std::mutex mu;
  static unsigned char *ptrs[64] = {nullptr};
  if (device < 0 || device >= 64) return ALNR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[device]) {
    void *p = nullptr;
    ALNR_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    ptrs[device] = static_cast<unsigned char *>(p);
  }
  *out = ptrs[device];
  return ALNR_OK;
}

}  // namespace alnr

extern "C" uint64_t alnr_noalloc_scratch_bytes(void) { return alnr::kNoAllocBytes; }

extern "C" int alnr_score_batch_noalloc(const alnr_params *p, const uint8_t *d_q, const uint64_t *d_q_off,
                                        const uint8_t *d_t, const uint64_t *d_t_off, uint64_t count,
                                        uint32_t max_query_len, uint32_t max_target_len, double *d_f,
                                        double *d_corner, uint32_t *d_end_yx, int32_t *d_status, void *stream) {
  using namespace alnr;
  if (!p || !p->matrix) return ALNR_INVALID_ARGUMENT;
  if (p->mode != ALNR_GLOBAL && p->mode != ALNR_LOCAL) return ALNR_INVALID_ARGUMENT;
  if (p->volume < 1 || p->volume > 32) return ALNR_MATRIX_SHAPE_ERROR;
  if (count == 0) return ALNR_OK;
  if (!d_q || !d_t || !d_q_off || !d_t_off || count >= (1ull << 31)) return ALNR_INVALID_ARGUMENT;
  if (max_query_len < 1 || max_target_len < 1) return ALNR_INVALID_ARGUMENT;
  // numeric domain as in batch.cu::classify: integers when every value the reference computes is one
  auto integral = [](double v) { return std::isfinite(v) && v == std::nearbyint(v) && std::fabs(v) <= 1048576.0; };
  const int VV = p->volume * p->volume;
  bool all_int = integral(p->del) && integral(p->ext);
  double mag = std::max(std::fabs(p->del), std::fabs(p->ext));
  for (int i = 0; i < VV; ++i) { all_int = all_int && integral(p->matrix[i]); mag = std::max(mag, std::fabs(p->matrix[i])); }
  if (p->mode == ALNR_GLOBAL && p->del == 0.0) all_int = false;  // -(x as f64) * 0.0 is -0.0 (simple/mod.rs:60)
  const double bound = ((double)max_query_len + (double)max_target_len + 2.0) * std::max(mag, 1.0);
  int numeric = (all_int && bound < 1073741824.0) ? ALNR_NUM_I32 : ALNR_NUM_F64;
  if (p->numeric == ALNR_NUM_F64) numeric = ALNR_NUM_F64;
  else if (p->numeric == ALNR_NUM_I16X2 || (p->numeric == ALNR_NUM_I32 && numeric != ALNR_NUM_I32)) {
    set_error_detail("alnr_score_batch_noalloc: forced numeric domain cannot represent this problem exactly");
    return ALNR_INVALID_ARGUMENT;
  }
  // the limit: one score column of max_target_len + 1 values per pair, 32 pairs per group
  const uint64_t elem = numeric == ALNR_NUM_F64 ? 8 : 4;
  const uint64_t need = (count + 31) / 32 * 32 * ((uint64_t)max_target_len + 1) * elem;
  if (need > kNoAllocBytes) {
    char buf[256];
    std::snprintf(buf, sizeof(buf),
                  "alnr_score_batch_noalloc: %llu pairs x (max_target_len %u + 1) x %llu B = %llu bytes of score "
                  "columns exceed the static scratch of %llu bytes (ALNR_NOALLOC_SCRATCH_BYTES); split the batch",
                  (unsigned long long)count, max_target_len, (unsigned long long)elem, (unsigned long long)need,
                  (unsigned long long)kNoAllocBytes);
    set_error_detail(buf);
    return ALNR_INVALID_ARGUMENT;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    (void)cudaGetLastError();
    set_error_detail("no usable CUDA device (this library has no CPU fallback)");
    return ALNR_CUDA_ERROR;
  }
  const int device = p->device < 0 ? 0 : p->device;
  if (device >= ndev) return ALNR_INVALID_ARGUMENT;
  ALNR_CUDA(cudaSetDevice(device));
  NoAllocArgs a;
  int st = noalloc_scratch_ptr(device, &a.scratch);
  if (st) return st;
  a.q = d_q; a.t = d_t; a.q_off = d_q_off; a.t_off = d_t_off;
  a.count = (uint32_t)count; a.n_cap = max_query_len; a.m_cap = max_target_len;
  a.mode = p->mode; a.V = p->volume; a.del = p->del; a.ext = p->ext;
  a.f = d_f; a.corner = d_corner; a.end_yx = d_end_yx; a.status = d_status;
  for (int i = 0; i < VV; ++i) a.matrix[i] = p->matrix[i];
  const unsigned blocks = (unsigned)((count + 127) / 128);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (numeric == ALNR_NUM_I32) score_exact_static_kernel<int><<<blocks, 128, (size_t)VV * 4, s>>>(a);
  else score_exact_static_kernel<double><<<blocks, 128, (size_t)VV * 8, s>>>(a);
  ALNR_LAUNCHED();
  return ALNR_OK;
}
