// kernels_strip.cu — host side of the s16x2 strip kernel family: eligibility, one-time packing,
// and the per-chunk launch sequence (speculative pass, re-run passes, exact leftovers).
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "kernels_exact.cuh"
#include "kernels_strip.cuh"
#include "strip_launch.h"

namespace alnr {

static const int kRedoPasses = 3;

// Packed 16-bit lanes are exact when: DNA-sized alphabet (4 score bytes per query symbol),
// scores fit a signed byte, 1 <= ext <= del (penalty update by one multiply-add without
// carries between lanes), every |H| < 2^15 with headroom, uniform lengths, and the strip's
// boundary row fits shared memory.
bool strip_eligible(const alnr_batch_impl &B, double bound) {
  if (B.V > 4 || !B.uniform || B.count < 1) return false;
  if (B.want_matrices) return false;  // single-pair calls that also return H and D
  if (B.n_max < 1 || B.m_max < 1) return false;
  if (bound >= 30000.0) return false;
  if (!(B.ext >= 1.0 && B.del >= B.ext && B.del <= 4096.0)) return false;
  for (double v : B.matrix) if (std::fabs(v) > 127.0) return false;
  if (((size_t)B.n_max + 1) * kStripBlock * 4 > 200 * 1024) return false;
  return true;
}

static uint32_t pick_rows(uint32_t m) {
  const uint32_t p16 = (m + 15) / 16 * 16 - m, p15 = (m + 14) / 15 * 15 - m;
  return p15 < p16 ? 15 : 16;
}

uint32_t d_rows_per_word(const alnr_batch_impl &B, uint32_t m) {
  return B.numeric == ALNR_NUM_I16X2 ? pick_rows(m) : 16;
}

// Incremental re-run (rerun_strip0_kernel): local mode with del != ext and at least two strips.
static bool strip_incremental(const alnr_batch_impl &B) {
  static const bool off = getenv("ALNR_NO_INCREMENTAL") != nullptr;
  return !off && B.mode == ALNR_LOCAL && B.del != B.ext && B.m_max > pick_rows(B.m_max);
}

int strip_prepare(alnr_batch_impl &B) {
  const uint32_t nwq = (B.n_max + 15) / 16, nwt = (B.m_max + 15) / 16;
  const uint64_t groups = (B.count + 31) / 32;
  ALNR_CUDA(B.d_pack.alloc(groups * 32 * ((uint64_t)nwq + nwt)));
  // speculation plumbing: two z buffers, two lists, counters (sized for the largest chunk)
  uint32_t cmax = 0;
  for (const Chunk &c : B.chunks) cmax = std::max(cmax, c.count);
  const uint32_t nzw = (B.n_max + 31) / 32;
  // layout of d_rerun: [counters 8 per chunk] then one block per chunk parity:
  // [list0 cmax][list1 cmax][z0 cmax*nzw][z1 cmax*nzw]
  B.rerun_words = 2ull * cmax + 2ull * cmax * nzw;
  ALNR_CUDA(B.d_rerun.alloc(8 * B.chunks.size() + (B.chunks.size() > 1 ? 2 : 1) * B.rerun_words));
  return ALNR_OK;
}

template <int R>
static int launch_strip(alnr_batch_impl &B, cudaStream_t stream, const StripArgs &a_in, uint32_t duos, bool local, bool stateful) {
  const size_t smem = a_in.hglobal ? 0 : (size_t)a_in.n * kStripBlock * 4;
  const unsigned vblocks = (duos + kStripBlock - 1) / kStripBlock;
  if (vblocks == 0) return ALNR_OK;
  StripArgs a = a_in;
  a.vblocks = vblocks;
#define ALNR_STRIP_LAUNCH(L, S, SC)                                                                   \
  do {                                                                                            \
    static bool attr_done = false;                                                                \
    static unsigned resident = 0;                                                                 \
    if (!attr_done) {                                                                             \
      ALNR_CUDA(cudaFuncSetAttribute(fill_strip_kernel<R, L, S, SC>,                              \
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));   \
      int occ = 0, sms = 0, dev = 0;                                                              \
      ALNR_CUDA(cudaGetDevice(&dev));                                                             \
      ALNR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));               \
      ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fill_strip_kernel<R, L, S, SC>,\
                                                              kStripBlock, smem));                \
      resident = (unsigned)std::max(1, occ) * (unsigned)std::max(1, sms);                         \
      attr_done = true;                                                                           \
    }                                                                                             \
    const unsigned blocks = a.hglobal ? std::min(vblocks, resident) : vblocks;                    \
    fill_strip_kernel<R, L, S, SC><<<blocks, kStripBlock, smem, stream>>>(a);                     \
  } while (0)
  if (B.score_only) {
    if (!local) ALNR_STRIP_LAUNCH(false, false, true);
    else if (stateful) ALNR_STRIP_LAUNCH(true, true, true);
    else ALNR_STRIP_LAUNCH(true, false, true);
  } else {
    if (!local) ALNR_STRIP_LAUNCH(false, false, false);
    else if (stateful) ALNR_STRIP_LAUNCH(true, true, false);
    else ALNR_STRIP_LAUNCH(true, false, false);
  }
#undef ALNR_STRIP_LAUNCH
  ALNR_LAUNCHED();
  return ALNR_OK;
}

// Pairs one resident wave of the strip kernel processes (SMs x resident blocks x 128): chunk
// sizes are multiples of it so every persistent block walks the same number of virtual blocks.
uint32_t strip_wave_pairs() {
  static uint32_t cached = 0;
  if (cached) return cached;
  int occ = 0, sms = 0, dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fill_strip_kernel<15, true, true>, kStripBlock, 0) !=
          cudaSuccess) {
    (void)cudaGetLastError();
    return 148u * 8u * 2u * kStripBlock;
  }
  cached = (uint32_t)std::max(1, occ) * (uint32_t)std::max(1, sms) * 2u * kStripBlock;
  return cached;
}

// Words of ONE boundary-row buffer of the strip kernels: a row of n words per physical thread of the
// largest launch (a resident wave, or fewer for small chunks).  The incremental re-run keeps two.
uint64_t strip_row_words(const alnr_batch_impl &B) {
  uint32_t cmax = 0;
  for (const Chunk &k : B.chunks) cmax = std::max(cmax, k.count);
  const uint64_t threads = std::min<uint64_t>(((uint64_t)cmax + 127) / 128 * 64, strip_wave_pairs() / 2) + kStripBlock;
  return (threads + 31) / 32 * 32 * (uint64_t)std::max(B.n_max, 1u);
}

namespace {
struct StripPlan {
  StripArgs a;
  uint32_t *counters, *lists[2], *zbuf[2];
  bool local, stateful;
  uint32_t R;
};

StripPlan strip_plan(alnr_batch_impl &B, const Chunk &c, const ChunkView &cv, const DLayout &dl, const FillOut &fo) {
  StripPlan P;
  const uint32_t nwq = (B.n_max + 15) / 16;
  const uint64_t groups = (B.count + 31) / 32;
  uint32_t cmax = 0;
  for (const Chunk &k : B.chunks) cmax = std::max(cmax, k.count);
  const uint32_t nzw = (B.n_max + 31) / 32;
  const size_t ci = &c - B.chunks.data();
  uint32_t *base = B.d_rerun.p + 8 * B.chunks.size() + (size_t)B.par * B.rerun_words;
  P.counters = B.d_rerun.p + 8 * ci;
  P.lists[0] = base; P.lists[1] = base + cmax;
  P.zbuf[0] = base + 2ull * cmax; P.zbuf[1] = base + 2ull * cmax + (uint64_t)cmax * nzw;
  P.local = B.mode == ALNR_LOCAL;
  P.stateful = P.local && B.del != B.ext;
  StripArgs &a = P.a;
  a.qpk = B.d_pack.p;
  a.tpk = B.d_pack.p + groups * 32 * nwq;
  a.nwq = nwq; a.nwt = (B.m_max + 15) / 16; a.n = B.n_max; a.m = B.m_max;
  a.first = c.first; a.count = c.count;
  a.order = cv.order;
  for (int qs = 0; qs < 4; ++qs) {
    uint32_t w = 0;
    for (int ts = 0; ts < 4; ++ts) {
      int v = (qs < B.V && ts < B.V) ? (int)B.matrix[(size_t)ts * B.V + qs] : 0;
      w |= ((uint32_t)v & 0xffu) << (8 * ts);
    }
    a.wtab[qs] = w;
  }
  a.del = (int)B.del; a.ext = (int)B.ext;
  a.k_one = 1u; a.k_mone = 0xffffffffu; a.k_bias = kBias; a.k_one2 = 0x00010001u;
  {
    double top = 0.0;
    for (double v : B.matrix) top = std::max(top, v);
    a.safe_mask = 0;
    static const bool off = getenv("ALNR_NO_SAFE_MASK") != nullptr;
    for (int ts = 0; ts < B.V && !off; ++ts)
      for (int qs = 0; qs < B.V; ++qs)
        if (B.matrix[(size_t)ts * B.V + qs] > top - B.ext) a.safe_mask |= 1u << (ts * 4 + qs);
  }
  a.D = B.Dp(); a.dl = dl; a.fo = fo;
  a.list = nullptr; a.list_count = nullptr; a.z_in = nullptr; a.z_out = P.zbuf[0]; a.nzw = nzw;
  a.fail_list = P.lists[0]; a.fail_count = P.counters + 0;
  static const bool hrow_global = getenv("ALNR_STRIP_HROW_SMEM") == nullptr;
  a.hglobal = hrow_global ? (uint32_t *)B.hcol_i() : nullptr;
  // the incremental re-run keeps a second row per resident thread, half a row buffer further on
  a.incremental = (strip_incremental(B) && a.hglobal) ? 1u : 0u;
  a.hglobal2 = strip_row_words(B);
  P.R = dl.rows_per_word;
  return P;
}
}  // namespace

int strip_main_pass(alnr_batch_impl &B, const Chunk &c, const ChunkView &cv, const DLayout &dl,
                    const FillOut &fo) {
  StripPlan P = strip_plan(B, c, cv, dl, fo);
  cudaStream_t stream = B.stream;
  // pack this chunk's sequences (2 bits per symbol, coalesced layout)
  if (B.packed_in) {  // compact input format: the words are there, only their order changes
    const uint64_t lg = (c.count + 31) / 32;
    uint64_t threads = lg * 32 * P.a.nwq;
    interleave2bit_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(
        B.d_q2raw.p, c.first, c.count, P.a.nwq, B.n_max, const_cast<uint32_t *>(P.a.qpk));
    ALNR_LAUNCHED();
    threads = lg * 32 * P.a.nwt;
    interleave2bit_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(
        B.d_t2raw.p, c.first, c.count, P.a.nwt, B.m_max, const_cast<uint32_t *>(P.a.tpk));
    ALNR_LAUNCHED();
  } else {
    const uint64_t lg = (c.count + 31) / 32;
    uint64_t threads = lg * 32 * P.a.nwq;
    pack2bit_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(
        B.d_q.p, B.d_qoff.p, cv.order, c.first, c.count, P.a.nwq, const_cast<uint32_t *>(P.a.qpk));
    ALNR_LAUNCHED();
    threads = lg * 32 * P.a.nwt;
    pack2bit_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(
        B.d_t.p, B.d_toff.p, cv.order, c.first, c.count, P.a.nwt, const_cast<uint32_t *>(P.a.tpk));
    ALNR_LAUNCHED();
  }
  if (P.stateful) ALNR_CUDA(cudaMemsetAsync(P.counters, 0, 8 * sizeof(uint32_t), stream));
  const uint32_t duos = (c.count + 63) / 64 * 32;
  return P.R == 15 ? launch_strip<15>(B, stream, P.a, duos, P.local, P.stateful)
                   : launch_strip<16>(B, stream, P.a, duos, P.local, P.stateful);
}

int strip_rerun_passes(alnr_batch_impl &B, const Chunk &c, const ChunkView &cv, const DLayout &dl,
                       const FillOut &fo) {
  StripPlan P = strip_plan(B, c, cv, dl, fo);
  cudaStream_t stream = B.s2;
  // Incremental pass first (rerun_strip0_kernel): strip 0 only, with the observed bits; what it
  // cannot settle goes on list 1.  The full passes then read list `off + pass`.
  const int off = (P.a.incremental && c.m_max > P.R) ? 1 : 0;
  if (off) {
    StripArgs r = P.a;
    r.list = P.lists[0]; r.list_count = P.counters + 0;
    r.z_in = P.zbuf[0]; r.z_out = nullptr;
    r.fail_list = P.lists[1]; r.fail_count = P.counters + 1;
    static unsigned resident = 0;
    if (!resident) {
      int occ = 0, sms = 0, dev = 0;
      ALNR_CUDA(cudaGetDevice(&dev));
      ALNR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      ALNR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rerun_strip0_kernel<15, false>, kStripBlock, 0));
      resident = (unsigned)std::max(1, occ) * (unsigned)std::max(1, sms);
    }
    // the list length is only known on the device: one resident wave of persistent blocks (the
    // physical-thread row buffer is sized for that), blocks past the list exit at once
    const unsigned blocks = std::min<unsigned>(resident, ((c.count + 1) / 2 + kStripBlock - 1) / kStripBlock);
    if (P.R == 15) {
      if (B.score_only) rerun_strip0_kernel<15, true><<<blocks, kStripBlock, 0, stream>>>(r);
      else rerun_strip0_kernel<15, false><<<blocks, kStripBlock, 0, stream>>>(r);
    } else {
      if (B.score_only) rerun_strip0_kernel<16, true><<<blocks, kStripBlock, 0, stream>>>(r);
      else rerun_strip0_kernel<16, false><<<blocks, kStripBlock, 0, stream>>>(r);
    }
    ALNR_LAUNCHED();
  }
  // re-run passes over the (device-side) fail list; grids are sized for the worst case and
  // exit at once past the list length, so no host synchronisation is needed.
  for (int pass = 0; pass < kRedoPasses; ++pass) {
    StripArgs r = P.a;
    const int li = pass + off;
    r.list = P.lists[li & 1]; r.list_count = P.counters + li;
    r.z_in = P.zbuf[pass & 1]; r.z_out = P.zbuf[(pass + 1) & 1];
    r.fail_list = P.lists[(li + 1) & 1]; r.fail_count = P.counters + li + 1;
    // the fail fraction shrinks quickly: size each grid for a generous bound, the kernel exits
    // past the list length (a list longer than the grid would be wrong, so bound by the chunk)
    const uint32_t duos = (c.count + 1) / 2;
    int st = P.R == 15 ? launch_strip<15>(B, stream, r, duos, true, true)
                       : launch_strip<16>(B, stream, r, duos, true, true);
    if (st) return st;
  }
  // leftovers: exact column-major kernel over the last list
  ExactArgs<int> e;
  e.cv = cv; e.mode = B.mode; e.V = B.V; e.del = (int)B.del; e.ext = (int)B.ext;
  e.matrix = B.d_mati.p; e.hcol = B.hcol_i(); e.hrows = c.m_max + 1; e.D = B.Dp(); e.dl = dl; e.fo = fo;
  e.H_out = nullptr; e.D_out = nullptr;
  e.list = P.lists[(kRedoPasses + off) & 1]; e.list_count = P.counters + kRedoPasses + off;
  fill_exact_kernel<int><<<(c.count + 127) / 128, 128, (size_t)B.V * B.V * sizeof(int), stream>>>(e);
  ALNR_LAUNCHED();
  return ALNR_OK;
}

// pairs re-run after pass k (k = 0: after the speculative pass), summed over chunks (diagnostic)
int strip_fail_counts(alnr_batch_impl &B, uint64_t out[4]) {
  std::vector<uint32_t> h(8 * B.chunks.size());
  ALNR_CUDA(cudaMemcpyAsync(h.data(), B.d_rerun.p, h.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, B.stream));
  ALNR_CUDA(cudaStreamSynchronize(B.stream));
  for (int k = 0; k < 4; ++k) out[k] = 0;
  for (size_t c = 0; c < B.chunks.size(); ++c)
    for (int k = 0; k < 4; ++k) out[k] += h[8 * c + k];
  return ALNR_OK;
}

}  // namespace alnr
