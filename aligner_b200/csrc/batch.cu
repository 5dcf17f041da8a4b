// batch.cu — host layer: packs a batch into device-resident buffers, picks the numeric
// domain, plans chunks against a scratch budget and enqueues fill + traceback kernels.
#include "batch.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <mutex>
#include <thread>
#include <type_traits>

#include "kernels_exact.cuh"
#include "kernels_trace.cuh"
#include "strip_launch.h"
#include "warp_args.h"
#include "wave_launch.h"

namespace alnr {

std::atomic<uint64_t> g_launches{0};
static thread_local std::string t_detail;
void set_error_detail(const std::string &s) { t_detail = s; }
const char *get_error_detail() { return t_detail.c_str(); }

// The two streams and three events of a call are kept for the next call on the same device: the
// reference's callers align pair after pair (statistics/mod.rs:270-278), and creating / destroying
// them was ~0.1 ms of a 0.36 ms single-pair call.  Work queued on a returned stream has been
// waited for by the call that queued it (run / download synchronise), so reuse is in order.
namespace {
struct StreamSet { int device; cudaStream_t s1, s2; cudaEvent_t e0, e1, e2; };
std::mutex g_sets_mutex;
std::vector<StreamSet> g_sets;  // at most kMaxSets idle sets
constexpr size_t kMaxSets = 32;
}  // namespace

static bool take_stream_set(int device, StreamSet &out) {
  std::lock_guard<std::mutex> lk(g_sets_mutex);
  for (size_t i = 0; i < g_sets.size(); ++i)
    if (g_sets[i].device == device) {
      out = g_sets[i];
      g_sets[i] = g_sets.back();
      g_sets.pop_back();
      return true;
    }
  return false;
}

alnr_batch_impl::~alnr_batch_impl() {
  // Early returns (a failed init / run) get here with work possibly still queued: the device buffers
  // go back to the pool and the stream set to the next caller only after both streams have drained.
  if (stream || stream2) {
    (void)cudaSetDevice(device);
    if (stream) (void)cudaStreamSynchronize(stream);
    if (stream2) (void)cudaStreamSynchronize(stream2);
    (void)cudaGetLastError();
  }
  for (cudaEvent_t e : fill_events) cudaEventDestroy(e);
  for (cudaEvent_t e : ev_p1) cudaEventDestroy(e);
  for (cudaEvent_t e : ev_p2) cudaEventDestroy(e);
  if (stream && stream2 && ev0 && ev1 && ev2) {
    std::lock_guard<std::mutex> lk(g_sets_mutex);
    if (g_sets.size() < kMaxSets) {
      g_sets.push_back(StreamSet{device, stream, stream2, ev0, ev1, ev2});
      return;
    }
  }
  if (ev0) cudaEventDestroy(ev0);
  if (ev1) cudaEventDestroy(ev1);
  if (ev2) cudaEventDestroy(ev2);
  if (stream2) cudaStreamDestroy(stream2);
  if (stream) cudaStreamDestroy(stream);
}

int alnr_batch_impl::fill_mark(bool) {
  if (!profile) return ALNR_OK;
  if (fill_events_used == fill_events.size()) {
    cudaEvent_t e;
    ALNR_CUDA(cudaEventCreate(&e));
    fill_events.push_back(e);
  }
  ALNR_CUDA(cudaEventRecord(fill_events[fill_events_used++], stream));
  return ALNR_OK;
}

int alnr_batch_impl::collect_fill_ms() {
  last_fill_ms = 0.f;
  for (size_t i = 0; i + 1 < fill_events_used; i += 2) {
    float t = 0.f;
    ALNR_CUDA(cudaEventElapsedTime(&t, fill_events[i], fill_events[i + 1]));
    last_fill_ms += t;
  }
  fill_events_used = 0;
  return ALNR_OK;
}

// Integer kernels are bit-exact iff every value the reference computes is an integer that
// the chosen lane width holds: |H| <= (n + m + 2) * max(|s|, |del|, |ext|).
int alnr_batch_impl::classify() {
  auto integral = [](double v) { return std::isfinite(v) && v == std::nearbyint(v) && std::fabs(v) <= 1048576.0; };
  bool all_int = integral(del) && integral(ext);
  double mag = std::max(std::fabs(del), std::fabs(ext));
  for (double v : matrix) { all_int = all_int && integral(v); mag = std::max(mag, std::fabs(v)); }
  // -(x as f64) * del with del == 0 is -0.0 in the reference (simple/mod.rs:60): keep f64 there.
  if (mode == ALNR_GLOBAL && del == 0.0) all_int = false;
  const double bound = ((double)n_max + (double)m_max + 2.0) * std::max(mag, 1.0);
  int best;
  if (!all_int || bound >= 1073741824.0) best = ALNR_NUM_F64;
  else if (strip_eligible(*this, bound)) best = ALNR_NUM_I16X2;
  else best = ALNR_NUM_I32;
  if (numeric == ALNR_NUM_AUTO) {
    numeric = best;
    // The packed strip kernel runs two pairs per THREAD: below ~16 k pairs it cannot fill the device
    // and one warp per pair is faster (measured on 150 bp reads: 8 k pairs 4.0 against 4.3 ms, 32 k
    // pairs 8.8 against 8.0 ms; a single 600 x 600 DNA pair 5.5 ms on one thread).  A forced
    // ALNR_NUM_I16X2 is still honoured (tests).
    if (best == ALNR_NUM_I16X2 && count < 16384) numeric = ALNR_NUM_I32;
    return ALNR_OK;
  }
  // a forced domain must still be exact
  if (numeric == ALNR_NUM_F64) return ALNR_OK;
  if (numeric == ALNR_NUM_I32 && best != ALNR_NUM_F64) return ALNR_OK;
  if (numeric == ALNR_NUM_I16X2 && best == ALNR_NUM_I16X2) return ALNR_OK;
  set_error_detail("forced numeric domain cannot represent this problem exactly");
  return ALNR_INVALID_ARGUMENT;
}


static __global__ void uniform_offsets_kernel(uint64_t *qo, uint64_t *to, uint64_t n_off, uint64_t n, uint64_t m) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_off) { qo[i] = i * n; to[i] = i * m; }
}

static uint64_t scratch_budget_bytes(double scale) {
  // cudaMemGetInfo was seen to take 0.1 - 60 ms per call: ask at most every few seconds per device
  struct Cached { uint64_t free_b = 0; std::chrono::steady_clock::time_point at; bool valid = false; };
  static thread_local Cached cache[16];
  int dev = 0;
  (void)cudaGetDevice(&dev);
  Cached &cd = cache[dev & 15];
  const auto now = std::chrono::steady_clock::now();
  size_t free_b = 0, total_b = 0;
  if (cd.valid && now - cd.at < std::chrono::seconds(5)) free_b = cd.free_b;
  else {
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return 1ull << 30;
    cd.free_b = free_b; cd.at = now; cd.valid = true;
  }
  // blocks the library's own pool holds are available to this call (pool_alloc hands them out or
  // gives them back to the driver on demand); `scale` halves after an out-of-memory (other host
  // threads plan against the same free bytes at the same time)
  uint64_t b = (uint64_t)((double)((free_b + pool_cached_bytes(dev)) / 3) * scale);
  const uint64_t cap = 24ull << 30;
  return std::max<uint64_t>(std::min<uint64_t>(b, cap), 64ull << 20);
}

int alnr_batch_impl::plan_chunks(const uint64_t *q_off, const uint64_t *t_off) {
  chunks.clear();
  if (count == 0) return ALNR_OK;
  std::vector<uint32_t> order;
  if (!uniform) {
    order.resize(count);
    std::iota(order.begin(), order.end(), 0u);
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
      uint64_t ma = t_off[a + 1] - t_off[a], mb = t_off[b + 1] - t_off[b];
      if (ma != mb) return ma > mb;
      uint64_t na = q_off[a + 1] - q_off[a], nb = q_off[b + 1] - q_off[b];
      if (na != nb) return na > nb;
      return a < b;
    });
    ALNR_CUDA(d_order.alloc(count));
    ALNR_CUDA(cudaMemcpyAsync(d_order.p, order.data(), count * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
    ALNR_CUDA(cudaStreamSynchronize(stream));
  }
  auto dims = [&](uint64_t slot, uint32_t &m, uint32_t &n) {
    uint64_t p = uniform ? slot : order[slot];
    m = (uint32_t)(t_off[p + 1] - t_off[p]);
    n = (uint32_t)(q_off[p + 1] - q_off[p]);
  };
  // suffix maximum of n over slots (m is non-increasing after the sort)
  std::vector<uint32_t> nsuf;
  if (!uniform) {
    nsuf.resize(count);
    uint32_t mx = 0;
    for (uint64_t s = count; s-- > 0;) { uint32_t m, n; dims(s, m, n); mx = std::max(mx, n); nsuf[s] = mx; }
  }
  const uint64_t budget = scratch_budget_bytes(budget_scale) / 2;  // two parity copies of the chunk scratch
  const uint64_t kStripWavePairs = numeric == ALNR_NUM_I16X2 ? strip_wave_pairs() : 1;
  const size_t num_sz = numeric == ALNR_NUM_F64 ? 8 : 4;
  uint64_t s = 0;
  while (s < count) {
    uint32_t m, n;
    dims(s, m, n);
    if (!uniform) n = nsuf[s];
    m = std::max(m, 1u); n = std::max(n, 1u);
    const DLayout pl = layout_dims(m, n, false);
    const uint64_t per_group = (score_only ? 0 : pl.words_for(32) * 4) + (((uint64_t)std::max(m, n) + 1) * num_sz + (uint64_t)(n + 1) * 4) * 32;
    uint64_t groups = std::max<uint64_t>(budget / per_group, 1);
    uint64_t c = std::min<uint64_t>(groups * 32, count - s);
    if (c < count - s) c = std::max<uint64_t>(c / 64 * 64, 32);  // keep duo-kernel warps whole
    if (c > max_chunk) c = max_chunk;
    // whole waves of the (persistent) strip kernel: every resident block then walks the same
    // number of virtual blocks and no SM idles at the end of a chunk
    if (numeric == ALNR_NUM_I16X2 && c < count - s && c >= kStripWavePairs) c = c / kStripWavePairs * kStripWavePairs;
    if (ramp && numeric == ALNR_NUM_I16X2 && count > 4 * kStripWavePairs) {
      // streaming call: a one-wave first chunk starts the compute after ~1 ms of copies; the last
      // chunks shrink 3, 2, 1 waves.  A chunk's results cross PCIe in ~1.1 ms per wave while the next
      // chunk computes ~1.7 ms per wave, so a chunk at least 0.65x its predecessor leaves no copy
      // backlog and only the last wave's ~1 ms of device->host copies follows the last kernel.
      const uint64_t W = kStripWavePairs, rem = count - s;
      if (s == 0) c = std::min<uint64_t>(c, W);
      else if (rem > 7 * W) c = std::min<uint64_t>(c, (rem - 6 * W) / 64 * 64);
      else if (rem > 3 * W + W / 2) c = std::min<uint64_t>(c, (rem - 3 * W) / 64 * 64);
      else if (rem > W + W / 2) c = std::min<uint64_t>(c, (rem - W) / 64 * 64);
    }
    uint32_t m1, n1;
    dims(s, m1, n1);
    bool wave = wave_eligible(*this, m1, n1);
    // the checkpointed re-fill has no stateful instantiation: such a pair stays on the warp kernel
    if (wave && mode == ALNR_LOCAL && del != ext && !score_only && wave_ckpt_group(m1, n1, budget)) wave = false;
    if (wave) { c = 1; m = m1; n = n1; }
    Chunk ck{(uint32_t)s, (uint32_t)c, m, n, 0, 0, 0, 0};
    if (wave) {
      const uint64_t pidx = uniform ? s : order[s];
      ck.wave = true; ck.wave_pair = (uint32_t)pidx; ck.q0 = q_off[pidx]; ck.t0 = t_off[pidx];
      ck.q1 = q_off[pidx + 1]; ck.t1 = t_off[pidx + 1];
      // a direction matrix larger than the scratch budget: keep band rows, re-fill group by group
      if (!score_only) ck.ckpt_group = wave_ckpt_group(m, n, budget);
    } else
    if (uniform) { ck.q0 = q_off[s]; ck.q1 = q_off[s + c]; ck.t0 = t_off[s]; ck.t1 = t_off[s + c]; }
    chunks.push_back(ck);
    s += c;
  }
  return ALNR_OK;
}

// One warp per pair pays off once a lane has real work (>= 24 target rows) or the batch is too
// small to fill the GPU with one thread per pair; tiny pairs in huge batches stay per-thread.
bool alnr_batch_impl::use_warp(uint32_t m) const {
  if (numeric == ALNR_NUM_I16X2) return false;
  static const bool off = getenv("ALNR_NO_WARP_KERNEL") != nullptr;
  if (off) return false;
  if (numeric == ALNR_NUM_F64) {
    // the FP64 warp kernel takes max by compare-select: needs finite inputs far from overflow
    auto ok = [](double v) { return std::fabs(v) < 1e290; };
    bool fin = ok(del) && ok(ext);
    for (double v : matrix) fin = fin && ok(v);
    if (!fin) return false;
  }
  return want_matrices || m >= 24 || count <= 4096;
}

DLayout alnr_batch_impl::layout_dims(uint32_t m, uint32_t n, bool wave) const {
  if (wave) return DLayout::make_blocked(wave_band_rows(m), m, n);
  if (use_warp(m)) return DLayout::make_skewed(want_matrices ? 8 : warp_pick_rows(m), m, n);
  // strip kernel: interleaved map (a warp's column store is one 128-byte line), bit-plane words
  return DLayout::make(d_rows_per_word(*this, m), m, n, false, numeric == ALNR_NUM_I16X2 ? 1 : 0);
}

DLayout alnr_batch_impl::layout_for(const Chunk &c) const {
  if (c.ckpt_group) return DLayout::make_blocked(4, std::min<uint64_t>(c.m_max, (uint64_t)c.ckpt_group * 128), c.n_max);
  return layout_dims(c.m_max, c.n_max, c.wave);
}

// Compact alignment output: identity pair order (uniform batch, so a chunk is a contiguous
// region of q_aln / t_aln) and move strings short enough for the two-phase traceback.
bool alnr_batch_impl::chunk_compact(const Chunk &c) const {
  const uint32_t nmw = (c.n_max + c.m_max + 1 + 15) / 16;
  return uniform && !c.wave && nmw <= 256 && c.count >= 32;
}

int alnr_batch_impl::init(const alnr_params *p, const alnr_batch *b, bool want_mats) {
  if (!p || !b || !p->matrix) return ALNR_INVALID_ARGUMENT;
  // ALNR_TIMING: host milliseconds of the stages of init()
  static const bool timing = getenv("ALNR_TIMING") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  std::string t_log;
  auto lap = [&](const char *what) {
    if (!timing) return;
    const auto now = std::chrono::steady_clock::now();
    char buf[64];
    std::snprintf(buf, sizeof(buf), " %s %.2f", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
    t_log += buf;
    t_prev = now;
  };
  if (p->mode != ALNR_GLOBAL && p->mode != ALNR_LOCAL) return ALNR_INVALID_ARGUMENT;
  if (p->volume < 1 || p->volume > 32) return ALNR_MATRIX_SHAPE_ERROR;
  if (b->count >= (1ull << 31)) return ALNR_INVALID_ARGUMENT;
  if (b->count && (!b->q_off || !b->t_off)) return ALNR_INVALID_ARGUMENT;
  if ((b->q2 != nullptr) != (b->t2 != nullptr)) return ALNR_INVALID_ARGUMENT;
  packed_in = b->q2 != nullptr;
  if (packed_in && p->volume != 4) {
    set_error_detail("2-bit packed input needs volume == 4 (DNA)");
    return ALNR_INVALID_ARGUMENT;
  }
  if (b->count && !packed_in && (!b->q || !b->t)) return ALNR_INVALID_ARGUMENT;
  if (packed_in && want_mats) return ALNR_INVALID_ARGUMENT;
  mode = p->mode; V = p->volume; del = p->del; ext = p->ext; numeric = p->numeric;
  device = p->device < 0 ? 0 : p->device;
  want_matrices = want_mats;
  matrix.assign(p->matrix, p->matrix + (size_t)V * V);
  count = b->count;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || device >= ndev) {
    (void)cudaGetLastError();
    set_error_detail("no usable CUDA device (this library has no CPU fallback)");
    return ALNR_CUDA_ERROR;
  }
  ALNR_CUDA(cudaSetDevice(device));
  StreamSet ss;
  if (take_stream_set(device, ss)) {
    stream = ss.s1; stream2 = ss.s2; ev0 = ss.e0; ev1 = ss.e1; ev2 = ss.e2;
  } else {
    ALNR_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    ALNR_CUDA(cudaStreamCreateWithFlags(&stream2, cudaStreamNonBlocking));
    ALNR_CUDA(cudaEventCreate(&ev0));
    ALNR_CUDA(cudaEventCreate(&ev1));
    ALNR_CUDA(cudaEventCreate(&ev2));
  }

  q_total = count ? b->q_off[count] : 0;
  t_total = count ? b->t_off[count] : 0;
  n_max = m_max = 0;
  uniform = true;
  {
    // offsets sweep (monotone, < 2^31 per sequence, uniform or ragged): 10^7 pairs are 160 MB of
    // offsets, so it runs on a few host threads
    // (memory-bound: up to 16 threads, as many as the host has)
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned nt = count >= (1u << 20) ? std::min(16u, hw) : 1u;
    struct Part { uint32_t n_max = 0, m_max = 0; bool uniform = true, bad = false; uint64_t mvw = 0; };
    std::vector<Part> parts(nt);
    const uint64_t n0 = count ? b->q_off[1] - b->q_off[0] : 0, m0 = count ? b->t_off[1] - b->t_off[0] : 0;
    auto sweep = [&](unsigned k) {
      Part &pt = parts[k];
      const uint64_t lo = count * k / nt, hi = count * (k + 1) / nt;
      uint64_t nmx = 0, mmx = 0, diff = 0, neg = 0, mvw = 0;
      for (uint64_t i = lo; i < hi; ++i) {
        const uint64_t n = b->q_off[i + 1] - b->q_off[i], m = b->t_off[i + 1] - b->t_off[i];
        neg |= (n | m) >> 31;  // covers offsets that go backwards (the difference wraps)
        mvw += (n + m + 15) >> 4;
        diff |= (n ^ n0) | (m ^ m0);
        nmx = n > nmx ? n : nmx;
        mmx = m > mmx ? m : mmx;
      }
      pt.bad = neg != 0; pt.uniform = diff == 0; pt.n_max = (uint32_t)nmx; pt.m_max = (uint32_t)mmx;
      pt.mvw = mvw;
    };
    if (nt == 1) sweep(0);
    else {
      std::vector<std::thread> th;
      for (unsigned k = 0; k < nt; ++k) th.emplace_back(sweep, k);
      for (auto &t : th) t.join();
    }
    mv_capacity = 0;
    for (const Part &pt : parts) {
      if (pt.bad) return ALNR_INVALID_ARGUMENT;
      mv_capacity += pt.mvw;
      uniform = uniform && pt.uniform;
      n_max = std::max(n_max, pt.n_max);
      m_max = std::max(m_max, pt.m_max);
    }
  }
  lap("streams+sweep");
  int st = classify();
  if (st) return st;
  if (packed_in) {
    // sequence i starts at word sum_{j<i} ceil(len_j / 16); uniform lengths need no table
    const uint64_t nwq = (n_max + 15) / 16, nwt = (m_max + 15) / 16;
    q2_words = count * nwq; t2_words = count * nwt;
    h_qwoff.clear(); h_twoff.clear();
    if (!uniform) {
      h_qwoff.resize(count + 1); h_twoff.resize(count + 1);
      uint64_t wq = 0, wt = 0;
      for (uint64_t i = 0; i < count; ++i) {
        h_qwoff[i] = wq; h_twoff[i] = wt;
        wq += (b->q_off[i + 1] - b->q_off[i] + 15) >> 4;
        wt += (b->t_off[i + 1] - b->t_off[i] + 15) >> 4;
      }
      h_qwoff[count] = q2_words = wq; h_twoff[count] = t2_words = wt;
    }
  }
  // byte codes on the device: not needed when the packed words feed the strip kernel directly and
  // no strings are expanded
  const bool need_bytes = !(packed_in && moves_out && numeric == ALNR_NUM_I16X2);
  ALNR_CUDA(d_q.alloc(need_bytes ? q_total + 16 : 16));
  ALNR_CUDA(d_t.alloc(need_bytes ? t_total + 16 : 16));
  ALNR_CUDA(d_qoff.alloc(count + 1));
  ALNR_CUDA(d_toff.alloc(count + 1));
  if (packed_in) {
    ALNR_CUDA(d_q2raw.alloc(q2_words + 4));
    ALNR_CUDA(d_t2raw.alloc(t2_words + 4));
    if (!uniform) {
      ALNR_CUDA(d_qwoff.alloc(count + 1));
      ALNR_CUDA(d_twoff.alloc(count + 1));
      ALNR_CUDA(cudaMemcpyAsync(d_qwoff.p, h_qwoff.data(), (count + 1) * 8, cudaMemcpyHostToDevice, stream));
      ALNR_CUDA(cudaMemcpyAsync(d_twoff.p, h_twoff.data(), (count + 1) * 8, cudaMemcpyHostToDevice, stream));
    }
  }
  if (count) {
    if (resident && packed_in) {
      ALNR_CUDA(cudaMemcpyAsync(d_q2raw.p, b->q2, q2_words * 4, cudaMemcpyHostToDevice, stream));
      ALNR_CUDA(cudaMemcpyAsync(d_t2raw.p, b->t2, t2_words * 4, cudaMemcpyHostToDevice, stream));
    } else if (resident) {
      ALNR_CUDA(cudaMemcpyAsync(d_q.p, b->q, q_total, cudaMemcpyHostToDevice, stream));
      ALNR_CUDA(cudaMemcpyAsync(d_t.p, b->t, t_total, cudaMemcpyHostToDevice, stream));
    }
    if (uniform && count >= 4096 && b->q_off[0] == 0 && b->t_off[0] == 0) {
      // equal lengths: the offsets are i * n and i * m — written on the device instead of
      // copying 16 bytes per pair over PCIe
      uniform_offsets_kernel<<<(unsigned)((count + 1 + 255) / 256), 256, 0, stream>>>(
          d_qoff.p, d_toff.p, count + 1, b->q_off[1], b->t_off[1]);
      ALNR_LAUNCHED();
    } else {
      ALNR_CUDA(cudaMemcpyAsync(d_qoff.p, b->q_off, (count + 1) * 8, cudaMemcpyHostToDevice, stream));
      ALNR_CUDA(cudaMemcpyAsync(d_toff.p, b->t_off, (count + 1) * 8, cudaMemcpyHostToDevice, stream));
    }
  }
  ALNR_CUDA(d_matd.alloc((size_t)V * V));
  ALNR_CUDA(d_mati.alloc((size_t)V * V));
  std::vector<int> mi((size_t)V * V);
  for (size_t i = 0; i < mi.size(); ++i) mi[i] = numeric == ALNR_NUM_F64 ? 0 : (int)matrix[i];
  ALNR_CUDA(cudaMemcpyAsync(d_matd.p, matrix.data(), matrix.size() * 8, cudaMemcpyHostToDevice, stream));
  ALNR_CUDA(cudaMemcpyAsync(d_mati.p, mi.data(), mi.size() * 4, cudaMemcpyHostToDevice, stream));
  ALNR_CUDA(cudaStreamSynchronize(stream));

  lap("inputs");
  // Scratch planned against the free memory; when an allocation still fails (several host threads
  // planning at once, fragmentation) the cached blocks are returned and the plan is redone with half
  // the budget — smaller chunks instead of ALNR_OUT_OF_MEMORY.
  for (int attempt = 0;; ++attempt) {
    st = plan_and_alloc(b);
    if (st != ALNR_OUT_OF_MEMORY || attempt == 4) break;
    pool_release_all();
    budget_scale *= 0.5;
  }
  if (timing) std::fprintf(stderr, "[alnr] init:%s ms\n", t_log.c_str());
  return st;
}

int alnr_batch_impl::plan_and_alloc(const alnr_batch *b) {
  for (cudaEvent_t e : ev_p1) cudaEventDestroy(e);
  for (cudaEvent_t e : ev_p2) cudaEventDestroy(e);
  ev_p1.clear(); ev_p2.clear();
  int st = plan_chunks(b->q_off, b->t_off);
  if (st) return st;

  ALNR_CUDA(d_f.alloc(count));
  ALNR_CUDA(d_corner.alloc(count));
  ALNR_CUDA(d_start.alloc(count * 2));
  ALNR_CUDA(d_status.alloc(count));
  ALNR_CUDA(d_coords.alloc(count * 4));
  ALNR_CUDA(d_alnoff.alloc(count));
  ALNR_CUDA(d_alnlen.alloc(count));
  if (score_only) moves_out = false;
  moves_direct = moves_out && !chunks.empty();
  for (const Chunk &c : chunks) moves_direct = moves_direct && chunk_compact(c);
  if (moves_out) {
    ALNR_CUDA(d_mvout.alloc(mv_capacity + 4));
    ALNR_CUDA(d_mvoff.alloc(count));
    ALNR_CUDA(d_mvbase.alloc(chunks.size() + 2));
  }
  const bool no_strings = score_only || moves_direct;
  ALNR_CUDA(d_qaln.alloc(no_strings ? 1 : q_total + t_total + count));
  ALNR_CUDA(d_taln.alloc(no_strings ? 1 : q_total + t_total + count));
  ALNR_CUDA(d_flags.alloc(4));
  ALNR_CUDA(d_rerun.alloc(count + 1));
  {
    uint32_t cmax = 0;
    for (const Chunk &c : chunks) cmax = std::max(cmax, c.count);
    // scan descriptors: a chunk's tiles, or (compact results derived from strings) the whole batch's
    ALNR_CUDA(d_desc.alloc((size_t)((moves_out && !moves_direct) ? count : cmax) / (kScanBlock * kScanItems) + 4));
    ALNR_CUDA(d_ctotal.alloc(chunks.size() + 1));
    uint64_t mwords = 1;
    for (const Chunk &c : chunks)
      if (chunk_compact(c) && !score_only)
        mwords = std::max<uint64_t>(mwords, (uint64_t)((c.count + 31) / 32 * 32) * ((c.n_max + c.m_max + 1 + 15) / 16));
      else if (c.wave && !score_only)  // long pair: move words, (cy, cx) per word, 3 tail words
        mwords = std::max<uint64_t>(mwords, 2 * (uint64_t)((c.n_max + c.m_max + 1 + 15) / 16) + 8);
    ALNR_CUDA(d_moves.alloc(2 * mwords));  // move words, then the start cell of every word
  }

  uint64_t dwords = 1, hcol = 1;
  for (const Chunk &c : chunks) {
    const DLayout dl = layout_for(c);
    if (!score_only) dwords = std::max<uint64_t>(dwords, dl.words_for(c.count));
    hcol = std::max<uint64_t>(hcol, (uint64_t)((c.count + 31) / 32) * 32 * ((uint64_t)std::max(c.m_max, c.n_max) + 1));
  }
  const size_t copies = chunks.size() > 1 ? 2 : 1;
  if (numeric == ALNR_NUM_I16X2) hcol = std::max<uint64_t>(hcol, 2 * strip_row_words(*this));
  d_words = (dwords + 3) / 4 * 4; hcol_elems = (hcol + 3) / 4 * 4;
  ALNR_CUDA(d_D.alloc(d_words * copies));
  if (numeric == ALNR_NUM_F64) ALNR_CUDA(d_hcol_d.alloc(hcol_elems * copies));
  else ALNR_CUDA(d_hcol_i.alloc(hcol_elems * copies));
  for (size_t i = 0; i < chunks.size(); ++i) {
    cudaEvent_t e1, e2;
    ALNR_CUDA(cudaEventCreateWithFlags(&e1, cudaEventDisableTiming));
    ev_p1.push_back(e1);
    ALNR_CUDA(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
    ev_p2.push_back(e2);
  }
  if (want_matrices && count == 1) {
    ALNR_CUDA(d_Hout.alloc(((size_t)m_max + 1) * ((size_t)n_max + 1)));
    ALNR_CUDA(d_Dout.alloc(((size_t)m_max + 1) * ((size_t)n_max + 1)));
  }
  if (numeric == ALNR_NUM_I16X2) {
    st = strip_prepare(*this);
    if (st) return st;
  } else if (mode == ALNR_LOCAL && del != ext) {
    uint32_t cmax = 0;
    for (const Chunk &c : chunks) cmax = std::max(cmax, c.count);
    ALNR_CUDA(d_wz.alloc(warp_scratch_words(cmax, (n_max + 31) / 32, true)));
  }
  return ALNR_OK;
}

template <typename T>
static int launch_exact(alnr_batch_impl &B, const Chunk &c, const ChunkView &cv, const DLayout &dl,
                        const FillOut &fo, T *hcol, const T *mat) {
  ExactArgs<T> a;
  a.cv = cv; a.mode = B.mode; a.V = B.V; a.del = (T)B.del; a.ext = (T)B.ext;
  a.matrix = mat; a.hcol = hcol; a.hrows = c.m_max + 1; a.D = B.Dp(); a.dl = dl; a.fo = fo;
  a.H_out = B.d_Hout.p; a.D_out = B.d_Dout.p;
  const uint32_t blocks = (c.count + 127) / 128;
  fill_exact_kernel<T><<<blocks, 128, (size_t)B.V * B.V * sizeof(T), B.stream>>>(a);
  ALNR_LAUNCHED();
  return ALNR_OK;
}

template <typename T>
static int launch_warp(alnr_batch_impl &B, const Chunk &c, const ChunkView &cv, const DLayout &dl,
                       const FillOut &fo, T *hcol, const T *mat) {
  WarpArgs<T> a;
  a.cv = cv; a.V = B.V; a.del = (T)B.del; a.ext = (T)B.ext;
  const bool local = B.mode == ALNR_LOCAL;
  a.stateful = local && B.del != B.ext;
  a.matrix = mat; a.item_matrices = nullptr;
  a.D = B.Dp(); a.dl = dl; a.fo = fo;
  a.rowbuf = hcol; a.rowbuf_stride = c.n_max + 1;
  a.nzw = (B.n_max + 31) / 32;
  uint32_t *counters = B.d_wz.p;
  static const bool no_inc = getenv("ALNR_NO_INCREMENTAL") != nullptr;
  warp_scratch_bind(a, counters, c.count, !no_inc);
  a.H_out = B.d_Hout.p; a.D_out = B.d_Dout.p;
  const bool mats = B.want_matrices && B.count == 1;
  if (a.stateful) ALNR_CUDA(cudaMemsetAsync(counters, 0, 8 * sizeof(uint32_t), B.stream));
  int st;
  if constexpr (std::is_same<T, int>::value) st = warp_launch_i32(B.stream, a, local, mats);
  else st = warp_launch_f64(B.stream, a, local, mats);
  if (st || !a.stateful) return st;
  // pairs whose speculation did not settle: exact column-major kernel over the fail list
  ExactArgs<T> e;
  e.cv = cv; e.mode = B.mode; e.V = B.V; e.del = (T)B.del; e.ext = (T)B.ext;
  e.matrix = mat; e.hcol = hcol; e.hrows = c.m_max + 1; e.D = B.Dp(); e.dl = dl; e.fo = fo;
  e.H_out = mats ? B.d_Hout.p : nullptr; e.D_out = mats ? B.d_Dout.p : nullptr;
  e.list = a.fail_list; e.list_count = a.fail_count;
  fill_exact_kernel<T><<<(c.count + 127) / 128, 128, (size_t)B.V * B.V * sizeof(T), B.stream>>>(e);
  ALNR_LAUNCHED();
  return ALNR_OK;
}

int alnr_batch_impl::run() {
  ALNR_CUDA(cudaSetDevice(device));
  if (count == 0) return ALNR_OK;
  ALNR_CUDA(cudaMemsetAsync(d_flags.p, 0, 4 * sizeof(int), stream));
  if (moves_out) ALNR_CUDA(cudaMemsetAsync(d_mvbase.p, 0, (chunks.size() + 2) * 8, stream));
  if (!uniform && packed_in) {
    int st = unpack_all();  // 2-bit codes are in range by construction: no symbol sweep
    if (st) return st;
  } else if (!uniform) {  // sorted order: chunks are not contiguous symbol ranges, sweep everything first
    any_symbol_out_of_range_kernel<<<148 * 8, 256, 0, stream>>>(d_q.p, q_total, V, d_flags.p);
    ALNR_LAUNCHED();
    any_symbol_out_of_range_kernel<<<148 * 8, 256, 0, stream>>>(d_t.p, t_total, V, d_flags.p);
    ALNR_LAUNCHED();
  }
  rerun_total = 0;
  for (size_t ci = 0; ci < chunks.size(); ++ci) {
    int st = run_chunk(ci);
    if (st) return st;
  }
  if (moves_out && !moves_direct) return moves_from_strings();
  return ALNR_OK;
}

// Compact input: byte codes of one chunk's sequences (uniform batches: a chunk is a contiguous
// range of sequences and of packed words).  Skipped when nothing downstream reads bytes.
int alnr_batch_impl::unpack_chunk(const Chunk &c) {
  if (packed_in && moves_out && numeric == ALNR_NUM_I16X2) return ALNR_OK;
  const uint32_t nwq = (n_max + 15) / 16, nwt = (m_max + 15) / 16;
  if (nwq) {
    unpack2bit_kernel<<<(unsigned)(((uint64_t)c.count * nwq + 255) / 256), 256, 0, stream>>>(
        d_q2raw.p, nullptr, nwq, d_qoff.p, c.first, c.count, d_q.p);
    ALNR_LAUNCHED();
  }
  if (nwt) {
    unpack2bit_kernel<<<(unsigned)(((uint64_t)c.count * nwt + 255) / 256), 256, 0, stream>>>(
        d_t2raw.p, nullptr, nwt, d_toff.p, c.first, c.count, d_t.p);
    ALNR_LAUNCHED();
  }
  return ALNR_OK;
}

int alnr_batch_impl::unpack_all() {
  const unsigned blocks = (unsigned)((count * 32 + 255) / 256);
  unpack2bit_kernel<<<blocks, 256, 0, stream>>>(d_q2raw.p, d_qwoff.p, 0, d_qoff.p, 0, count, d_q.p);
  ALNR_LAUNCHED();
  unpack2bit_kernel<<<blocks, 256, 0, stream>>>(d_t2raw.p, d_twoff.p, 0, d_toff.p, 0, count, d_t.p);
  ALNR_LAUNCHED();
  return ALNR_OK;
}

// Compact results for batches whose chunks ran a string traceback (ragged batches, long pairs, tiny
// batches): one scan of the word counts in PAIR order over the whole batch, then the move words are
// read off the aligned strings.  On `stream`, after the last chunk's second phase.
int alnr_batch_impl::moves_from_strings() {
  const uint32_t tiles = (uint32_t)((count + kScanBlock * kScanItems - 1) / (kScanBlock * kScanItems));
  ALNR_CUDA(cudaMemsetAsync(d_desc.p, 0, ((size_t)tiles + 1) * 8, stream));
  scan_lengths_kernel<true><<<tiles, kScanBlock, 0, stream>>>(d_alnlen.p, d_mvoff.p, 0, (uint32_t)count, 0,
                                                              reinterpret_cast<unsigned long long *>(d_desc.p),
                                                              d_mvbase.p + chunks.size());
  ALNR_LAUNCHED();
  strings_to_moves_kernel<<<(unsigned)((count * 32 + 127) / 128), 128, 0, stream>>>(
      d_qaln.p, d_taln.p, d_alnoff.p, d_alnlen.p, d_mvoff.p, count, d_mvout.p);
  ALNR_LAUNCHED();
  return ALNR_OK;
}

// Phase 1 on `stream`, phase 2 on the second stream (on `stream` too when profiling the fill
// or when the batch is a single chunk).  Chunk k+2 reuses chunk k's parity copy of the scratch,
// so its phase 1 waits for phase 2 of chunk k; `stream` also waits for the LAST phase 2 so that
// the caller only ever has to synchronise `stream`.
int alnr_batch_impl::run_chunk(size_t ci) {
  static const bool serial_env = getenv("ALNR_SERIAL_PHASES") != nullptr;
  // a checkpointed pair re-fills from the band rows in its second phase: the next long pair's first
  // phase must not overwrite them, so that chunk stays on one stream
  const bool overlap = chunks.size() > 1 && !profile && !serial_env && !chunks[ci].ckpt_group;
  s2 = overlap ? stream2 : stream;
  par = chunks.size() > 1 ? (int)(ci & 1) : 0;
  if (overlap && ci >= 2) ALNR_CUDA(cudaStreamWaitEvent(stream, ev_p2[ci - 2], 0));
  int st = phase1(ci);
  if (st) return st;
  if (overlap) {
    ALNR_CUDA(cudaEventRecord(ev_p1[ci], stream));
    ALNR_CUDA(cudaStreamWaitEvent(stream2, ev_p1[ci], 0));
  }
  st = phase2(ci);
  if (st) return st;
  if (overlap) {
    ALNR_CUDA(cudaEventRecord(ev_p2[ci], stream2));
    if (ci + 1 == chunks.size()) ALNR_CUDA(cudaStreamWaitEvent(stream, ev_p2[ci], 0));
  }
  return ALNR_OK;
}

int alnr_batch_impl::phase1(size_t ci) {
  const Chunk &c = chunks[ci];
  FillOut fo{d_f.p, d_corner.p, d_start.p, d_status.p};
  ChunkView cv{d_q.p, d_t.p, d_qoff.p, d_toff.p, uniform ? nullptr : d_order.p, c.first, c.count};
  const DLayout dl = layout_for(c);
  if (packed_in && uniform) {
    int stu = unpack_chunk(c);
    if (stu) return stu;
  }
  // symbol validation: one coalesced sweep; the per-pair scan only runs if it found something
  // (2-bit packed input is in range by construction)
  if ((uniform || c.wave) && !packed_in) {
    const unsigned gq = (unsigned)std::min<uint64_t>(148 * 8, (c.q1 - c.q0) / 4096 + 1);
    const unsigned gt = (unsigned)std::min<uint64_t>(148 * 8, (c.t1 - c.t0) / 4096 + 1);
    any_symbol_out_of_range_kernel<<<gq, 256, 0, stream>>>(d_q.p + c.q0, c.q1 - c.q0, V, d_flags.p);
    ALNR_LAUNCHED();
    any_symbol_out_of_range_kernel<<<gt, 256, 0, stream>>>(d_t.p + c.t0, c.t1 - c.t0, V, d_flags.p);
    ALNR_LAUNCHED();
  }
  validate_kernel<<<(c.count + 127) / 128, 128, 0, stream>>>(cv, V, d_status.p, d_flags.p);
  ALNR_LAUNCHED();
  int st = fill_mark(false);
  if (st) return st;
  if (c.wave) st = wave_run_pair(*this, c, c.ckpt_group ? layout_dims(c.m_max, c.n_max, true) : dl, fo);
  else if (numeric == ALNR_NUM_F64 && dl.skewed) st = launch_warp<double>(*this, c, cv, dl, fo, hcol_d(), d_matd.p);
  else if (numeric == ALNR_NUM_I32 && dl.skewed) st = launch_warp<int>(*this, c, cv, dl, fo, hcol_i(), d_mati.p);
  else if (numeric == ALNR_NUM_F64) st = launch_exact<double>(*this, c, cv, dl, fo, hcol_d(), d_matd.p);
  else if (numeric == ALNR_NUM_I32) st = launch_exact<int>(*this, c, cv, dl, fo, hcol_i(), d_mati.p);
  else st = strip_main_pass(*this, c, cv, dl, fo);
  if (st) return st;
  if (numeric != ALNR_NUM_I16X2 || !(mode == ALNR_LOCAL && del != ext)) st = fill_mark(true);
  return st;
}

int alnr_batch_impl::phase2(size_t ci) {
  const Chunk &c = chunks[ci];
  FillOut fo{d_f.p, d_corner.p, d_start.p, d_status.p};
  TraceOut to{want_coords ? d_coords.p : nullptr, d_alnoff.p, d_alnlen.p, d_qaln.p, d_taln.p, nullptr};
  ChunkView cv{d_q.p, d_t.p, d_qoff.p, d_toff.p, uniform ? nullptr : d_order.p, c.first, c.count};
  const DLayout dl = layout_for(c);
  if (numeric == ALNR_NUM_I16X2 && !c.wave && mode == ALNR_LOCAL && del != ext) {
    int st = strip_rerun_passes(*this, c, cv, dl, fo);  // on s2
    if (st) return st;
    st = fill_mark(true);  // only recorded when profiling, i.e. when s2 == stream
    if (st) return st;
  }
  if (score_only) return ALNR_OK;  // f, corner and the arg-max cell are final after the fill
  TraceArgs ta{cv, mode, V, Dp(), dl, fo, to};
  const uint32_t nmw = (c.n_max + c.m_max + 1 + 15) / 16;  // 2-bit move words per pair
  if (chunk_compact(c) && moves_direct) {
    // compact result format: move strings (A), scan of the word counts continuing the running total
    // of the earlier chunks (S), dense move words in pair order (G) — no strings are expanded
    const uint32_t stride = (c.count + 31) / 32 * 32;
    const uint32_t tiles = (c.count + kScanBlock * kScanItems - 1) / (kScanBlock * kScanItems);
    trace_moves_kernel<<<(c.count + 127) / 128, 128, 0, s2>>>(ta, d_moves.p, nullptr, stride);
    ALNR_LAUNCHED();
    ALNR_CUDA(cudaMemsetAsync(d_desc.p, 0, ((size_t)tiles + 1) * 8, s2));
    scan_lengths_kernel<true><<<tiles, kScanBlock, 0, s2>>>(d_alnlen.p, d_mvoff.p, c.first, c.count, 0,
                                                            reinterpret_cast<unsigned long long *>(d_desc.p),
                                                            d_mvbase.p + ci + 1, d_mvbase.p + ci);
    ALNR_LAUNCHED();
    moves_gather_kernel<<<(c.count + 127) / 128, 128, 0, s2>>>(d_moves.p, stride, d_alnlen.p, d_mvoff.p, c.first,
                                                               c.count, d_mvout.p);
  } else if (chunk_compact(c)) {
    // alignments of the chunk back to back from the start of its region (kernels_trace.cuh A/S/B)
    const uint32_t stride = (c.count + 31) / 32 * 32;
    const uint32_t tiles = (c.count + kScanBlock * kScanItems - 1) / (kScanBlock * kScanItems);
    uint32_t *cxy = d_moves.p + (size_t)stride * nmw;
    trace_moves_kernel<<<(c.count + 127) / 128, 128, 0, s2>>>(ta, d_moves.p, cxy, stride);
    ALNR_LAUNCHED();
    ALNR_CUDA(cudaMemsetAsync(d_desc.p, 0, ((size_t)tiles + 1) * 8, s2));
    scan_lengths_kernel<false><<<tiles, kScanBlock, 0, s2>>>(d_alnlen.p, d_alnoff.p, c.first, c.count,
                                                      c.q0 + c.t0 + c.first,
                                                      reinterpret_cast<unsigned long long *>(d_desc.p),
                                                      d_ctotal.p + ci);
    ALNR_LAUNCHED();
    const uint32_t G = trace_expand_group(c.n_max, c.m_max);
    trace_expand_kernel<<<(c.count + G - 1) / G, kExpandThreads, trace_expand_smem(G, c.n_max, c.m_max), s2>>>(
        ta, d_moves.p, cxy, stride, G, c.n_max, c.m_max, nmw);
  } else if (c.ckpt_group) {
    return wave_ckpt_trace(*this, c, dl, ta, s2);
  } else if (dl.blocked == 4 && c.count == 1) {
    // long pair, 4 x 4 tiles: cooperative window walk that emits moves, then a parallel expansion
    LongTrace lt{d_moves.p, d_moves.p + nmw, d_moves.p + 3 * (size_t)nmw};
    trace_walk_coop_kernel<<<1, 32, 0, s2>>>(ta, lt);
    ALNR_LAUNCHED();
    trace_cells_kernel<<<1, 1024, 0, s2>>>(ta, lt);
    ALNR_LAUNCHED();
    trace_expand_long_kernel<<<(nmw + 255) / 256, 256, 0, s2>>>(ta, lt);
  } else if (dl.blocked) {
    traceback_warp_kernel<<<c.count, 32, 0, s2>>>(ta);
  } else if ((size_t)nmw * kTraceBlock * 4 <= 44 * 1024 && to.freq == nullptr) {
    traceback_coop_kernel<<<(c.count + kTraceBlock - 1) / kTraceBlock, kTraceBlock,
                            (size_t)nmw * kTraceBlock * 4, s2>>>(ta, nmw);
  } else {
    traceback_kernel<<<(c.count + 127) / 128, 128, 0, s2>>>(ta);
  }
  ALNR_LAUNCHED();
  return ALNR_OK;
}

// Pipelined host->host call: chunk k's inputs are copied on `s_in` while chunk k-1 computes
// on `stream` and chunk k-2's results drain on `s_out`.  Needs identity order (uniform
// lengths) so that a chunk is a contiguous range of every input and output array.
int alnr_batch_impl::stream_all(const alnr_batch *b, alnr_batch_out *out) {
  ALNR_CUDA(cudaSetDevice(device));
  if (count == 0) return ALNR_OK;
  cudaStream_t s_in = nullptr, s_out = nullptr;
  ALNR_CUDA(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
  ALNR_CUDA(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
  std::vector<cudaEvent_t> ev_in(chunks.size(), nullptr), ev_done(chunks.size(), nullptr);
  int st = ALNR_OK;
  auto fail = [&](cudaError_t e, const char *what) { st = cuda_fail(e, what, __FILE__, __LINE__); };
  for (size_t i = 0; i < chunks.size() && !st; ++i) {
    cudaError_t e = cudaEventCreateWithFlags(&ev_in[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev_done[i], cudaEventDisableTiming);
    if (e != cudaSuccess) fail(e, "cudaEventCreate");
  }
  if (!st) {
    cudaError_t e = cudaMemsetAsync(d_flags.p, 0, 4 * sizeof(int), stream);
    if (e == cudaSuccess && moves_out) e = cudaMemsetAsync(d_mvbase.p, 0, (chunks.size() + 2) * 8, stream);
    if (e != cudaSuccess) fail(e, "memset flags");
  }
  rerun_total = 0;
  // ALNR_TIMING: per-chunk timeline (H2D end, compute begin/end, D2H end) relative to the first copy
  static const bool trace = getenv("ALNR_TIMING") != nullptr;
  std::vector<cudaEvent_t> tev;
  auto mark = [&](cudaStream_t sx) {
    if (!trace) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, sx);
    tev.push_back(e);
  };
  mark(s_in);
  // chunk totals of the compact alignment output come back through pinned memory; the host waits
  // for chunk k's total only after chunk k+1 has been enqueued, so the device never idles
  uint64_t *h_tot = nullptr;
  bool h_tot_owned = true;
  if (!st) {
    // pinned words for the chunk totals: one small buffer per calling thread, kept (cudaMallocHost
    // takes about a millisecond, which is most of a small batch's call)
    struct PinnedWords {
      uint64_t *p = nullptr;
      ~PinnedWords() { if (p) { cudaFreeHost(p); (void)cudaGetLastError(); } }
    };
    static thread_local PinnedWords cached;
    constexpr size_t kCachedWords = 4096;
    cudaError_t e = cudaSuccess;
    if (chunks.size() + 1 <= kCachedWords) {
      if (!cached.p) e = cudaMallocHost(&cached.p, kCachedWords * sizeof(uint64_t));
      h_tot = cached.p;
      h_tot_owned = false;
    } else {
      e = cudaMallocHost(&h_tot, (chunks.size() + 1) * sizeof(uint64_t));
    }
    if (e != cudaSuccess) fail(e, "cudaMallocHost totals");
  }
  auto drain = [&](size_t ci) {  // device -> host copies of chunk ci on s_out
    const Chunk &c = chunks[ci];
    cudaError_t e = cudaSuccess;
    const size_t f0 = c.first, fc = c.count;
    const size_t a0 = c.q0 + c.t0 + f0;  // first byte of the chunk's alignment region
    size_t abytes = (c.q1 + c.t1 + f0 + fc) - a0;
    // s_out waits for THIS chunk only (a wait for the next chunk queued ahead of these copies
    // would hold them back by a whole chunk period)
    e = cudaStreamWaitEvent(s_out, ev_done[ci], 0);
    if (e == cudaSuccess && chunk_compact(c)) {
      e = cudaEventSynchronize(ev_done[ci]);  // h_tot[ci] was copied before ev_done was recorded
      if (e == cudaSuccess) abytes = (size_t)h_tot[ci];
    }
    auto cp = [&](void *dst, const void *src, size_t off, size_t bytes) {
      if (dst && bytes && e == cudaSuccess)
        e = cudaMemcpyAsync((char *)dst + off, (const char *)src + off, bytes, cudaMemcpyDeviceToHost, s_out);
    };
    cp(out->f, d_f.p, f0 * 8, fc * 8);
    cp(out->corner, d_corner.p, f0 * 8, fc * 8);
    cp(out->coords, d_coords.p, f0 * 32, fc * 32);
    cp(out->end_yx, d_start.p, f0 * 8, fc * 8);
    cp(out->aln_len, d_alnlen.p, f0 * 4, fc * 4);
    cp(out->status, d_status.p, f0 * 4, fc * 4);
    if (moves_out) {
      // h_tot[ci] = move words up to and including this chunk (running total kept on the device)
      const size_t w0 = ci ? (size_t)h_tot[ci - 1] : 0, w1 = (size_t)h_tot[ci];
      cp(out->moves_off, d_mvoff.p, f0 * 8, fc * 8);
      cp(out->moves, d_mvout.p, w0 * 4, (w1 - w0) * 4);
    } else {
      cp(out->aln_off, d_alnoff.p, f0 * 8, fc * 8);
      cp(out->q_aln, d_qaln.p, a0, abytes);
      cp(out->t_aln, d_taln.p, a0, abytes);
    }
    if (e != cudaSuccess) fail(e, "D2H chunk");
    mark(s_out);
  };
  // Host->device copies run kAhead chunks ahead of the compute.  Queuing all of them up front keeps
  // the link busy in the upward direction for the first half of the call, and the device->host
  // copies of the first chunks then run at half speed (measured: 22-33 GB/s next to a saturating
  // upload, 51 GB/s alone) and pile up a backlog that is still draining after the last kernel.
  const size_t kAhead = 3, kDrainLag = 2;
  auto upload = [&](size_t ci) {
    if (ci >= chunks.size() || st) return;
    const Chunk &c = chunks[ci];
    cudaError_t e;
    if (packed_in) {  // uniform lengths: the chunk's sequences are a contiguous range of packed words
      const size_t nwq = (n_max + 15) / 16, nwt = (m_max + 15) / 16;
      e = cudaMemcpyAsync(d_q2raw.p + c.first * nwq, b->q2 + c.first * nwq, c.count * nwq * 4, cudaMemcpyHostToDevice, s_in);
      if (e == cudaSuccess)
        e = cudaMemcpyAsync(d_t2raw.p + c.first * nwt, b->t2 + c.first * nwt, c.count * nwt * 4, cudaMemcpyHostToDevice, s_in);
    } else {
      e = cudaMemcpyAsync(d_q.p + c.q0, b->q + c.q0, c.q1 - c.q0, cudaMemcpyHostToDevice, s_in);
      if (e == cudaSuccess) e = cudaMemcpyAsync(d_t.p + c.t0, b->t + c.t0, c.t1 - c.t0, cudaMemcpyHostToDevice, s_in);
    }
    if (e == cudaSuccess) e = cudaEventRecord(ev_in[ci], s_in);
    if (e != cudaSuccess) fail(e, "H2D chunk");
  };
  for (size_t ci = 0; ci < kAhead; ++ci) upload(ci);
  for (size_t ci = 0; ci < chunks.size() && !st; ++ci) {
    const Chunk &c = chunks[ci];
    upload(ci + kAhead);
    if (st) break;
    cudaError_t e = cudaStreamWaitEvent(stream, ev_in[ci], 0);
    if (e != cudaSuccess) { fail(e, "H2D wait"); break; }
    mark(stream);
    st = run_chunk(ci);
    if (st) break;
    mark(s2);
    // s2 = the stream phase 2 of this chunk ran on: its results are complete there
    if (chunk_compact(c))
      e = cudaMemcpyAsync(h_tot + ci, moves_out ? d_mvbase.p + ci + 1 : d_ctotal.p + ci, sizeof(uint64_t),
                          cudaMemcpyDeviceToHost, s2);
    if (e == cudaSuccess) e = cudaEventRecord(ev_done[ci], s2);
    if (e != cudaSuccess) { fail(e, "chunk events"); break; }
    // drain() blocks the host until its chunk is complete: draining chunk ci-2 here keeps the
    // kernels of chunk ci queued behind chunk ci-1 on the device (draining ci-1 would enqueue every
    // chunk only when the one before it has finished its second phase)
    if (ci >= kDrainLag) drain(ci - kDrainLag);
  }
  for (size_t ci = chunks.size() > kDrainLag ? chunks.size() - kDrainLag : 0; ci < chunks.size() && !st; ++ci) drain(ci);
  cudaError_t e1 = cudaStreamSynchronize(s_in), e2 = cudaStreamSynchronize(stream), e3 = cudaStreamSynchronize(s_out);
  if (e2 == cudaSuccess) e2 = cudaStreamSynchronize(stream2);
  if (!st && e1 != cudaSuccess) fail(e1, "sync s_in");
  if (!st && e2 != cudaSuccess) fail(e2, "sync compute");
  if (!st && e3 != cudaSuccess) fail(e3, "sync s_out");
  if (trace && !st && tev.size() == 1 + 3 * chunks.size()) {
    // order of marks: [0] start, then per chunk ci: compute begin, compute end; the D2H end of
    // chunk ci - kDrainLag is recorded inside iteration ci (after its compute-end mark), the last ones at the end
    auto ms = [&](size_t k) { float t = 0; cudaEventElapsedTime(&t, tev[0], tev[k]); return t; };
    std::vector<size_t> c0(chunks.size()), c1(chunks.size()), d2h(chunks.size());
    size_t k = 1;
    for (size_t ci = 0; ci < chunks.size(); ++ci) {
      c0[ci] = k++; c1[ci] = k++;
      if (ci >= kDrainLag) d2h[ci - kDrainLag] = k++;
    }
    for (size_t ci = chunks.size() > kDrainLag ? chunks.size() - kDrainLag : 0; ci < chunks.size(); ++ci) d2h[ci] = k++;
    for (size_t ci = 0; ci < chunks.size(); ++ci)
      std::fprintf(stderr, "[alnr] chunk %2zu: compute %7.2f -> %7.2f  d2h_end %7.2f ms\n", ci,
                   ms(c0[ci]), ms(c1[ci]), ms(d2h[ci]));
  }
  for (cudaEvent_t e : tev) cudaEventDestroy(e);
  for (size_t i = 0; i < chunks.size(); ++i) {
    if (ev_in[i]) cudaEventDestroy(ev_in[i]);
    if (ev_done[i]) cudaEventDestroy(ev_done[i]);
  }
  cudaStreamDestroy(s_in);
  cudaStreamDestroy(s_out);
  if (h_tot && h_tot_owned) cudaFreeHost(h_tot);
  return st;
}

int alnr_batch_impl::download_scores(double *f, double *corner, uint32_t *end_yx, int32_t *status) {
  ALNR_CUDA(cudaSetDevice(device));
  if (count == 0) return ALNR_OK;
  auto cp = [&](void *dst, const void *src, size_t bytes) -> cudaError_t {
    if (!dst) return cudaSuccess;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream);
  };
  ALNR_CUDA(cp(f, d_f.p, count * 8));
  ALNR_CUDA(cp(corner, d_corner.p, count * 8));
  ALNR_CUDA(cp(end_yx, d_start.p, count * 8));
  ALNR_CUDA(cp(status, d_status.p, count * 4));
  ALNR_CUDA(cudaStreamSynchronize(stream));
  return ALNR_OK;
}

int alnr_batch_impl::download(alnr_batch_out *out) {
  if (!out) return ALNR_INVALID_ARGUMENT;
  if (score_only) {  // no alignments were produced
    if (out->q_aln || out->t_aln || out->coords || out->aln_off || out->aln_len) return ALNR_INVALID_ARGUMENT;
    return download_scores(out->f, out->corner, nullptr, out->status);
  }
  if (out->moves && !moves_out) {
    set_error_detail("compact results need a handle made by alnr_batch_create_compact");
    return ALNR_INVALID_ARGUMENT;
  }
  if (moves_out) return download_moves(out);
  ALNR_CUDA(cudaSetDevice(device));
  if (count == 0) return ALNR_OK;
  auto cp = [&](void *dst, const void *src, size_t bytes) -> cudaError_t {
    if (!dst) return cudaSuccess;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream);
  };
  ALNR_CUDA(cp(out->f, d_f.p, count * 8));
  ALNR_CUDA(cp(out->corner, d_corner.p, count * 8));
  ALNR_CUDA(cp(out->coords, d_coords.p, count * 32));
  ALNR_CUDA(cp(out->end_yx, d_start.p, count * 8));
  ALNR_CUDA(cp(out->aln_off, d_alnoff.p, count * 8));
  ALNR_CUDA(cp(out->aln_len, d_alnlen.p, count * 4));
  ALNR_CUDA(cp(out->status, d_status.p, count * 4));
  ALNR_CUDA(cp(out->q_aln, d_qaln.p, q_total + t_total + count));
  ALNR_CUDA(cp(out->t_aln, d_taln.p, q_total + t_total + count));
  ALNR_CUDA(cudaStreamSynchronize(stream));
  if (numeric == ALNR_NUM_I16X2 && mode == ALNR_LOCAL && del != ext) {
    uint64_t fc[4];
    int st = strip_fail_counts(*this, fc);
    if (st) return st;
    rerun_total = fc[0];
    static const bool timing = getenv("ALNR_TIMING") != nullptr;
    if (timing)  // pairs on the list after the speculative pass, the incremental pass, full passes 1 and 2
      std::fprintf(stderr, "[alnr] speculation lists: %llu %llu %llu %llu of %llu pairs\n", (unsigned long long)fc[0],
                   (unsigned long long)fc[1], (unsigned long long)fc[2], (unsigned long long)fc[3], (unsigned long long)count);
  }
  return ALNR_OK;
}

int alnr_batch_impl::download_moves(alnr_batch_out *out) {
  if (out->q_aln || out->t_aln || out->aln_off) {
    set_error_detail("compact results: q_aln / t_aln / aln_off must be NULL (alnr_expand_alignments makes them)");
    return ALNR_INVALID_ARGUMENT;
  }
  ALNR_CUDA(cudaSetDevice(device));
  if (count == 0) return ALNR_OK;
  uint64_t words = 0;  // running total after the last chunk (or after the whole-batch scan)
  ALNR_CUDA(cudaMemcpyAsync(&words, d_mvbase.p + chunks.size(), 8, cudaMemcpyDeviceToHost, stream));
  ALNR_CUDA(cudaStreamSynchronize(stream));
  auto cp = [&](void *dst, const void *src, size_t bytes) -> cudaError_t {
    if (!dst || !bytes) return cudaSuccess;
    return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream);
  };
  ALNR_CUDA(cp(out->f, d_f.p, count * 8));
  ALNR_CUDA(cp(out->corner, d_corner.p, count * 8));
  ALNR_CUDA(cp(out->coords, d_coords.p, count * 32));
  ALNR_CUDA(cp(out->end_yx, d_start.p, count * 8));
  ALNR_CUDA(cp(out->aln_len, d_alnlen.p, count * 4));
  ALNR_CUDA(cp(out->status, d_status.p, count * 4));
  ALNR_CUDA(cp(out->moves_off, d_mvoff.p, count * 8));
  ALNR_CUDA(cp(out->moves, d_mvout.p, words * 4));
  ALNR_CUDA(cudaStreamSynchronize(stream));
  return ALNR_OK;
}

}  // namespace alnr
