// kernels_trace.cuh — validation and on-device traceback.
#pragma once
#include "common.cuh"

namespace alnr {

// One thread per pair: empty sequences and out-of-alphabet symbols become statuses
// (the reference panics: simple/mod.rs:103-104 unwrap, matrix[[t,q]] bounds check).
// The per-symbol scan only runs when the coalesced sweep below flagged the batch.
static __global__ void validate_kernel(ChunkView cv, int V, int32_t *status, const int *flag) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= cv.count) return;
  const uint32_t pair = cv.order ? cv.order[cv.first + slot] : cv.first + slot;
  const uint32_t sp = cv.seq(pair);
  if (cv.pwm_len) {  // PWM window: an empty window is a valid (empty) alignment, pwm/mod.rs:76-81
    const uint8_t *qp, *tp;
    uint32_t n, m;
    cv.pair_dims(sp, qp, tp, n, m);
    int st = 0;
    if (*flag) for (uint32_t i = 0; i < m && !st; ++i) if (tp[i] >= 4) st = 12;
    status[pair] = st;
    return;
  }
  const uint64_t q0 = cv.q_off[sp], q1 = cv.q_off[sp + 1];
  const uint64_t t0 = cv.t_off[sp], t1 = cv.t_off[sp + 1];
  int st = 0;
  if (q1 == q0 || t1 == t0) st = 10;  // ALNR_EMPTY_INPUT
  if (*flag) {
    for (uint64_t i = q0; i < q1 && !st; ++i) if (cv.q[i] >= V) st = 12;  // SYMBOL_OUT_OF_RANGE
    for (uint64_t i = t0; i < t1 && !st; ++i) if (cv.t[i] >= V) st = 12;
  }
  status[pair] = st;
}

// Fast whole-buffer symbol check (coalesced 16-byte loads); flags[0] != 0 iff some byte >= V.
static __global__ void any_symbol_out_of_range_kernel(const uint8_t *buf, uint64_t len, int V, int *flag) {
  const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  bool bad = false;
  // unaligned head (chunk ranges start anywhere), 16-byte body, tail
  uint64_t head = (16 - ((uintptr_t)buf & 15)) & 15;
  if (head > len) head = len;
  for (uint64_t i = tid; i < head; i += stride) bad |= buf[i] >= V;
  const uint64_t nvec = (len - head) / 16;
  const uint4 *v = reinterpret_cast<const uint4 *>(buf + head);
  for (uint64_t i = tid; i < nvec; i += stride) {
    uint4 w = v[i];
    uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int b = 0; b < 4; ++b) bad |= ((ws[k] >> (8 * b)) & 0xff) >= (uint32_t)V;
  }
  for (uint64_t i = head + nvec * 16 + tid; i < len; i += stride) bad |= buf[i] >= V;
  if (bad) *flag = 1;
}

struct TraceArgs {
  ChunkView cv;
  int mode;
  int V;
  const uint32_t *D;
  DLayout dl;
  FillOut fo;
  TraceOut to;
};

// One thread per pair walks the 2-bit D scratch from the start cell until Beginning
// (simple/mod.rs:99-130 global, :213-245 local).  The reference seeds the output with the
// start pair BEFORE the walk and reverses at the end; writing backwards from the end of the
// pair's slot yields the same string right-aligned with no reversal pass.
static __global__ void __launch_bounds__(128) traceback_kernel(TraceArgs a) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= a.cv.count) return;
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first + slot] : a.cv.first + slot;
  const uint32_t sp = a.cv.seq(pair);
  const uint64_t qo = a.cv.q_off[sp], to = a.cv.t_off[sp];
  const uint32_t n = (uint32_t)(a.cv.q_off[sp + 1] - qo);
  const uint32_t m = (uint32_t)(a.cv.t_off[sp + 1] - to);
  const uint64_t slot_begin = qo + to + pair;
  const uint64_t slot_end = slot_begin + n + m + 1;
  double *freq = a.to.freq ? a.to.freq + (size_t)pair * a.V * a.V : nullptr;
  if (freq) for (int i = 0; i < a.V * a.V; ++i) freq[i] = 0.0;
  if (a.fo.status[pair] != 0) {
    if (a.to.aln_len) a.to.aln_len[pair] = 0;
    if (a.to.aln_off) a.to.aln_off[pair] = slot_end;
    if (a.to.coords) for (int k = 0; k < 4; ++k) a.to.coords[4 * (size_t)pair + k] = 0;
    return;
  }
  const uint8_t *q = a.cv.q + qo;
  const uint8_t *t = a.cv.t + to;
  const bool glob = a.mode == 0;
  const uint32_t sy = a.fo.start[2 * (size_t)pair], sx = a.fo.start[2 * (size_t)pair + 1];
  uint32_t cy = sy, cx = sx;
  uint64_t pos = slot_end;
  const uint32_t rpw = a.dl.rows_per_word;
  const bool wr = a.to.q_aln != nullptr && a.to.t_aln != nullptr;

  auto emit = [&](uint8_t qc, uint8_t tc) {
    --pos;
    if (wr) { a.to.q_aln[pos] = qc; a.to.t_aln[pos] = tc; }
    if (freq && qc != kBlank && tc != kBlank) freq[(int)tc * a.V + qc] += 1.0;  // alignment.rs:16-20
  };
  emit(q[cx - 1], t[cy - 1]);  // seed (simple/mod.rs:102-105, :213-216)
  for (;;) {
    int d;
    if (cy == 0 || cx == 0) {
      if (!glob || (cy == 0 && cx == 0)) d = DIR_BEGINNING;
      else d = (cy == 0) ? DIR_LEFT : DIR_TOP;  // simple/mod.rs:59-67
    } else {
      const uint32_t w = a.D[a.dl.word(slot, (cy - 1) / rpw, cx - 1)];
      d = a.dl.get(w, (cy - 1) % rpw, cx - 1);
    }
    if (d == DIR_BEGINNING) break;
    if (d == DIR_TOP) { emit(kBlank, t[cy - 1]); --cy; }
    else if (d == DIR_LEFT) { emit(q[cx - 1], kBlank); --cx; }
    else { emit(q[cx - 1], t[cy - 1]); --cx; --cy; }
  }
  if (a.to.aln_len) a.to.aln_len[pair] = (uint32_t)(slot_end - pos);
  if (a.to.aln_off) a.to.aln_off[pair] = pos;
  if (a.to.coords) {
    uint64_t *c = a.to.coords + 4 * (size_t)pair;
    if (glob) { c[0] = 1; c[1] = n; c[2] = 1; c[3] = m; }                      // simple/mod.rs:137
    else { c[0] = cx + 1; c[1] = sx + 1; c[2] = cy + 1; c[3] = sy + 1; }        // :254-257
  }
}

// Two-phase traceback for short pairs (n + m + 1 moves fit shared memory).
//  phase A: one thread per pair chases the 2-bit directions and records its moves
//           (2 bits each) in shared memory — no global stores on the dependent chain;
//  phase B: each warp expands the 32 move strings of its pairs cooperatively: 32 moves at a
//           time, symbol positions by ballot/popc prefix counts, 32 consecutive output bytes
//           per store instead of one scattered byte per thread.
constexpr int kTraceBlock = 128;

static __global__ void __launch_bounds__(kTraceBlock) traceback_coop_kernel(TraceArgs a, uint32_t nmw) {
  extern __shared__ uint32_t mv[];  // mv[w * kTraceBlock + tid]
  __shared__ uint32_t s_len[kTraceBlock], s_sy[kTraceBlock], s_sx[kTraceBlock];
  const uint32_t tid = threadIdx.x;
  const uint32_t slot = blockIdx.x * kTraceBlock + tid;
  const bool glob = a.mode == 0;
  const uint32_t rpw = a.dl.rows_per_word;
  uint32_t K = 0xffffffffu;  // marker: no alignment for this thread
  if (slot < a.cv.count) {
    const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first + slot] : a.cv.first + slot;
    const uint64_t qo = a.cv.q_off[pair], to = a.cv.t_off[pair];
    const uint32_t n = (uint32_t)(a.cv.q_off[pair + 1] - qo);
    const uint32_t m = (uint32_t)(a.cv.t_off[pair + 1] - to);
    const uint64_t slot_end = qo + to + pair + n + m + 1;
    if (a.fo.status[pair] != 0) {
      if (a.to.aln_len) a.to.aln_len[pair] = 0;
      if (a.to.aln_off) a.to.aln_off[pair] = slot_end;
      if (a.to.coords) for (int k = 0; k < 4; ++k) a.to.coords[4 * (size_t)pair + k] = 0;
    } else {
      const uint32_t sy = a.fo.start[2 * (size_t)pair], sx = a.fo.start[2 * (size_t)pair + 1];
      uint32_t cy = sy, cx = sx, acc = 0;
      K = 0;
      for (;;) {
        int d;
        if (cy == 0 || cx == 0) {
          if (!glob || (cy == 0 && cx == 0)) d = DIR_BEGINNING;
          else d = (cy == 0) ? DIR_LEFT : DIR_TOP;
        } else {
          const uint32_t w = a.D[a.dl.word(slot, (cy - 1) / rpw, cx - 1)];
          d = a.dl.get(w, (cy - 1) % rpw, cx - 1);
        }
        if (d == DIR_BEGINNING) break;
        acc |= (uint32_t)d << (2 * (K & 15));
        ++K;
        if ((K & 15) == 0) { mv[((K >> 4) - 1) * kTraceBlock + tid] = acc; acc = 0; }
        if (d != DIR_LEFT) --cy;
        if (d != DIR_TOP) --cx;
      }
      if (K & 15) mv[(K >> 4) * kTraceBlock + tid] = acc;
      if (a.to.aln_len) a.to.aln_len[pair] = K + 1;
      if (a.to.aln_off) a.to.aln_off[pair] = slot_end - (K + 1);
      if (a.to.coords) {
        uint64_t *c = a.to.coords + 4 * (size_t)pair;
        if (glob) { c[0] = 1; c[1] = n; c[2] = 1; c[3] = m; }
        else { c[0] = cx + 1; c[1] = sx + 1; c[2] = cy + 1; c[3] = sy + 1; }
      }
      s_sy[tid] = sy; s_sx[tid] = sx;
    }
  }
  s_len[tid] = K;
  __syncwarp();
  if (a.to.q_aln == nullptr || a.to.t_aln == nullptr) return;

  // ---- phase B: this warp expands its own 32 pairs ------------------------------------
  const uint32_t lane = tid & 31, wbase = tid & ~31u;
  const uint32_t ltmask = (1u << lane) - 1u;
  for (uint32_t pi = 0; pi < 32; ++pi) {
    const uint32_t col = wbase + pi;
    const uint32_t Kp = s_len[col];
    if (Kp == 0xffffffffu) continue;
    const uint32_t pslot = blockIdx.x * kTraceBlock + col;
    const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first + pslot] : a.cv.first + pslot;
    const uint64_t qo = a.cv.q_off[pair], to = a.cv.t_off[pair];
    const uint8_t *q = a.cv.q + qo, *t = a.cv.t + to;
    const uint64_t slot_end = qo + to + pair + (a.cv.q_off[pair + 1] - qo) + (a.cv.t_off[pair + 1] - to) + 1;
    uint32_t cy = s_sy[col], cx = s_sx[col];
    if (lane == 0) {  // seed column (simple/mod.rs:102-105, :213-216)
      a.to.q_aln[slot_end - 1] = q[cx - 1];
      a.to.t_aln[slot_end - 1] = t[cy - 1];
    }
    for (uint32_t i0 = 0; i0 < Kp; i0 += 32) {
      const uint32_t i = i0 + lane;
      const bool valid = i < Kp;
      const uint32_t code = valid ? (mv[(i >> 4) * kTraceBlock + col] >> (2 * (i & 15))) & 3u : 3u;
      const bool cq = valid && code != DIR_TOP, ct = valid && code != DIR_LEFT;
      const uint32_t bq = __ballot_sync(0xffffffffu, cq), bt = __ballot_sync(0xffffffffu, ct);
      if (valid) {
        const uint32_t xi = cx - __popc(bq & ltmask), yi = cy - __popc(bt & ltmask);
        const uint64_t pos = slot_end - 2 - i;
        a.to.q_aln[pos] = cq ? q[xi - 1] : (uint8_t)kBlank;
        a.to.t_aln[pos] = ct ? t[yi - 1] : (uint8_t)kBlank;
      }
      cx -= __popc(bq);
      cy -= __popc(bt);
    }
  }
}

}  // namespace alnr

namespace alnr {

// ---- compact output for batches of equal-length short pairs (identity pair order) -----------
// Three launches: (A) one thread per pair chases the directions and stores its 2-bit move string
// in global memory, (S) a single-pass chained scan turns the alignment lengths into offsets —
// alignment i starts at base + sum of aln_len of the chunk's earlier pairs — and (B) each block
// expands the move strings of 32 pairs into the packed output through shared memory.  The chunk's alignments are back to back, so
// the device->host copy moves only the bytes that exist.

// (A) mvg[w * stride + slot] = moves 16w .. 16w+15, cxyg[w * stride + slot] = cx | cy << 16 of the
// walk before move 16w (word 0: the start cell); writes aln_len (0 for rejected pairs) and coords.
// (A per-thread queue of speculatively fetched direction words along the diagonal was measured at
// 2.6 ms per launch against 0.9 ms for this plain walk: the warp's 32 walks already overlap their
// misses, the queue only added instructions and DRAM sectors.)
static __global__ void __launch_bounds__(128) trace_moves_kernel(TraceArgs a, uint32_t *mvg, uint32_t *cxyg,
                                                                 uint32_t stride) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= a.cv.count) return;
  const uint32_t pair = a.cv.first + slot;
  const bool glob = a.mode == 0;
  const uint32_t rpw = a.dl.rows_per_word;
  if (a.fo.status[pair] != 0) {
    a.to.aln_len[pair] = 0;
    if (a.to.coords) for (int k = 0; k < 4; ++k) a.to.coords[4 * (size_t)pair + k] = 0;
    return;
  }
  const uint32_t n = (uint32_t)(a.cv.q_off[pair + 1] - a.cv.q_off[pair]);
  const uint32_t m = (uint32_t)(a.cv.t_off[pair + 1] - a.cv.t_off[pair]);
  const uint32_t sy = a.fo.start[2 * (size_t)pair], sx = a.fo.start[2 * (size_t)pair + 1];
  uint32_t cy = sy, cx = sx, acc = 0, K = 0;
  uint32_t *mv = mvg + slot, *cw = cxyg + slot;
  for (;;) {
    if (cxyg && (K & 15) == 0) cw[(size_t)(K >> 4) * stride] = cx | (cy << 16);
    int d;
    if (cy == 0 || cx == 0) {
      if (!glob || (cy == 0 && cx == 0)) d = DIR_BEGINNING;
      else d = (cy == 0) ? DIR_LEFT : DIR_TOP;
    } else {
      const uint32_t w = a.D[a.dl.word(slot, (cy - 1) / rpw, cx - 1)];
      d = a.dl.get(w, (cy - 1) % rpw, cx - 1);
    }
    if (d == DIR_BEGINNING) break;
    acc |= (uint32_t)d << (2 * (K & 15));
    ++K;
    if ((K & 15) == 0) { mv[(size_t)((K >> 4) - 1) * stride] = acc; acc = 0; }
    if (d != DIR_LEFT) --cy;
    if (d != DIR_TOP) --cx;
  }
  if (K & 15) mv[(size_t)(K >> 4) * stride] = acc;
  a.to.aln_len[pair] = K + 1;
  if (a.to.coords) {
    uint64_t *c = a.to.coords + 4 * (size_t)pair;
    if (glob) { c[0] = 1; c[1] = n; c[2] = 1; c[3] = m; }
    else { c[0] = cx + 1; c[1] = sx + 1; c[2] = cy + 1; c[3] = sy + 1; }
  }
}

// (S) aln_off[first + i] = base + sum_{j<i} aln_len[first + j]; *total = sum of all.
// Single pass: tiles of 1024 pairs are taken in arrival order (ticket), every tile publishes a
// 64-bit descriptor (flag 1 = its own sum, 2 = inclusive prefix; low 62 bits = value) and warp 0
// looks back 32 descriptors at a time.  desc[0] is the ticket, desc[1..] the tiles; all zeroed.
// WORDS: the scanned quantity is the number of 2-bit move words of an alignment, ceil((len - 1) / 16)
// (compact result format), `base` comes from *base_ptr (the running total of the chunks before this
// one) and *total continues it.
constexpr int kScanBlock = 256, kScanItems = 4;

template <bool WORDS>
static __global__ void __launch_bounds__(kScanBlock) scan_lengths_kernel(const uint32_t *len, uint64_t *off,
                                                                         uint32_t first, uint32_t count, uint64_t base,
                                                                         unsigned long long *descbuf, uint64_t *total,
                                                                         const uint64_t *base_ptr = nullptr) {
  __shared__ uint32_t s_wsum[kScanBlock / 32], s_tile;
  __shared__ unsigned long long s_excl;
  const uint32_t tid = threadIdx.x, lane = tid & 31;
  if (tid == 0) s_tile = atomicAdd(reinterpret_cast<uint32_t *>(descbuf), 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  volatile unsigned long long *desc = descbuf + 1;
  const uint32_t i0 = (tile * kScanBlock + tid) * kScanItems;
  uint32_t v[kScanItems], tsum = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    v[k] = (i0 + k < count) ? len[first + i0 + k] : 0u;
    if (WORDS) v[k] = (v[k] + 14u) >> 4;  // len = moves + 1 (0: no alignment)
    tsum += v[k];
  }
  uint32_t inc = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= (uint32_t)o) inc += u;
  }
  if (lane == 31) s_wsum[tid >> 5] = inc;
  __syncthreads();
  uint32_t woff = 0, T = 0;
#pragma unroll
  for (int w = 0; w < kScanBlock / 32; ++w) { if ((uint32_t)w < (tid >> 5)) woff += s_wsum[w]; T += s_wsum[w]; }
  if (tid < 32) {
    const unsigned long long vmask = (1ull << 62) - 1;
    unsigned long long excl = 0;
    if (tile > 0) {
      if (tid == 0) desc[tile] = (1ull << 62) | T;
      int hi = (int)tile - 1;
      for (;;) {
        const int j = hi - (int)tid;
        unsigned long long d = 2ull << 62;  // before tile 0: an empty inclusive prefix
        if (j >= 0) { do { d = desc[j]; } while ((d >> 62) == 0); }
        const uint32_t pfx = __ballot_sync(0xffffffffu, (d >> 62) == 2);
        const int stop = pfx ? __ffs(pfx) - 1 : 31;
        unsigned long long s = ((int)tid <= stop) ? (d & vmask) : 0ull;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        excl += s;
        if (pfx) break;
        hi -= 32;
      }
    }
    if (tid == 0) {
      desc[tile] = (2ull << 62) | (excl + T);
      s_excl = excl;
      if (tile + 1 == gridDim.x) *total = (base_ptr ? *base_ptr : 0ull) + excl + T;
    }
  }
  __syncthreads();
  uint64_t o = base + (base_ptr ? *base_ptr : 0ull) + s_excl + woff + (inc - tsum);
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    if (i0 + k < count) off[first + i0 + k] = o;
    o += v[k];
  }
}

// (B) a block expands the move strings of G consecutive pairs.  Their symbols are one contiguous
// range of q / t and their alignments one contiguous range of q_aln / t_aln, so both pass through
// shared memory: the inputs arrive and the outputs leave as aligned 16-byte vectors (the shared
// image has the global range's offset within its first 16 bytes), and in between one thread per
// (pair, move word) walks its 16 moves from the word's start cell (cxyg) with byte accesses to
// shared memory only.  The reference seeds the string with the start pair and reverses it
// (simple/mod.rs:102-105,:129-130): move i lands at aln_off + aln_len - 2 - i, the seed at
// aln_off + aln_len - 1.
constexpr int kExpandThreads = 256;

// shared bytes of trace_expand_kernel for G pairs of n x m symbols
static inline size_t trace_expand_smem(uint32_t G, uint32_t n, uint32_t m) {
  auto r16 = [](size_t v) { return (v + 15) / 16 * 16; };
  return r16((size_t)G * 8) + r16((size_t)G * n + 32) + r16((size_t)G * m + 32) + 2 * r16((size_t)G * (n + m + 1) + 32);
}

// pairs per block: 32 (one warp reads a 128-byte line of move words) unless the strings are long
static inline uint32_t trace_expand_group(uint32_t n, uint32_t m) {
  uint32_t G = 32;
  while (G > 1 && trace_expand_smem(G, n, m) > 46 * 1024) G >>= 1;
  return G;
}

static __device__ __forceinline__ void stage_in(uint8_t *dst, const uint8_t *src, uint32_t bytes, uint32_t &shift) {
  // dst[shift + i] = src[i]: whole aligned vectors around the range (the buffers are padded)
  shift = (uint32_t)((uintptr_t)src & 15);
  const uint4 *v = reinterpret_cast<const uint4 *>(src - shift);
  for (uint32_t i = threadIdx.x * 16; i < shift + bytes; i += kExpandThreads * 16)
    *reinterpret_cast<uint4 *>(dst + i) = __ldg(v + (i >> 4));
}

static __device__ __forceinline__ void stage_out(uint8_t *dst, const uint8_t *src, uint32_t bytes, uint32_t shift) {
  // dst[i] = src[shift + i], dst - shift 16-byte aligned: vectors inside the range, bytes at its ends
  for (uint32_t i = threadIdx.x * 16; i < shift + bytes; i += kExpandThreads * 16) {
    if (i >= shift && i + 16 <= shift + bytes) {
      *reinterpret_cast<uint4 *>(dst - shift + i) = *reinterpret_cast<const uint4 *>(src + i);
    } else {
      const uint32_t lo = i > shift ? i : shift, hi = i + 16 < shift + bytes ? i + 16 : shift + bytes;
      for (uint32_t j = lo; j < hi; ++j) dst[(size_t)j - shift] = src[j];
    }
  }
}

static __global__ void __launch_bounds__(kExpandThreads) trace_expand_kernel(TraceArgs a, const uint32_t *mvg,
                                                                             const uint32_t *cxyg, uint32_t stride,
                                                                             uint32_t G, uint32_t n, uint32_t m,
                                                                             uint32_t nmw) {
  extern __shared__ __align__(16) uint8_t sm_expand[];
  auto r16 = [](uint32_t v) { return (v + 15u) / 16u * 16u; };
  const uint32_t slot0 = blockIdx.x * G;
  if (slot0 >= a.cv.count) return;
  const uint32_t cnt = min(G, a.cv.count - slot0);
  const uint32_t pair0 = a.cv.first + slot0;
  uint32_t *s_len = reinterpret_cast<uint32_t *>(sm_expand);  // [G] aln_len
  uint32_t *s_end = s_len + G;                                // [G] end of the alignment, relative to the block's range
  uint8_t *q_in = sm_expand + r16(G * 8);
  uint8_t *t_in = q_in + r16(G * n + 32);
  uint8_t *q_out = t_in + r16(G * m + 32);
  uint8_t *t_out = q_out + r16(G * (n + m + 1) + 32);
  const uint64_t out0 = a.to.aln_off[pair0];
  if (threadIdx.x < cnt) {
    const uint32_t L = a.to.aln_len[pair0 + threadIdx.x];
    s_len[threadIdx.x] = L;
    s_end[threadIdx.x] = (uint32_t)(a.to.aln_off[pair0 + threadIdx.x] - out0) + L;
  }
  uint32_t shq, sht;
  stage_in(q_in, a.cv.q + a.cv.q_off[pair0], cnt * n, shq);
  stage_in(t_in, a.cv.t + a.cv.t_off[pair0], cnt * m, sht);
  __syncthreads();
  const uint32_t total = s_end[cnt - 1];
  uint8_t *gq = a.to.q_aln + out0, *gt = a.to.t_aln + out0;
  const uint32_t oq = (uint32_t)((uintptr_t)gq & 15), ot = (uint32_t)((uintptr_t)gt & 15);
  for (uint32_t u = threadIdx.x; u < G * nmw; u += kExpandThreads) {
    const uint32_t g = G == 32 ? (u & 31u) : u % G, w = G == 32 ? (u >> 5) : u / G;
    if (g >= cnt) continue;
    const uint32_t L = s_len[g];
    if (L == 0 || (w != 0 && w * 16 >= L - 1)) continue;
    const uint32_t cxy = cxyg[(size_t)w * stride + slot0 + g];
    uint32_t cx = cxy & 0xffffu, cy = cxy >> 16;
    const uint8_t *qs = q_in + shq + g * n - 1, *ts = t_in + sht + g * m - 1;  // qs[x] = symbol x (1-based)
    const uint32_t e = s_end[g];
    if (w == 0) { q_out[oq + e - 1] = qs[cx]; t_out[ot + e - 1] = ts[cy]; }
    if (w * 16 >= L - 1) continue;
    uint32_t mvw = mvg[(size_t)w * stride + slot0 + g];
    const uint32_t steps = min(16u, L - 1 - w * 16);
    uint32_t pos = e - 2 - w * 16;
    for (uint32_t j = 0; j < steps; ++j, --pos, mvw >>= 2) {
      const uint32_t code = mvw & 3u;
      uint8_t bq = (uint8_t)kBlank, bt = (uint8_t)kBlank;
      if (code != DIR_TOP) { bq = qs[cx]; --cx; }
      if (code != DIR_LEFT) { bt = ts[cy]; --cy; }
      q_out[oq + pos] = bq;
      t_out[ot + pos] = bt;
    }
  }
  __syncthreads();
  stage_out(gq, q_out, total, oq);
  stage_out(gt, t_out, total, ot);
}

// ---- compact result format (alnr_batch_out.moves) ------------------------------------------------
// (G) after (A) and the word scan: the move words of every pair, [w * stride + slot] in the scratch,
// become one dense run per pair at mvoff[pair].  A warp takes 32 slots; lane <-> slot, so the reads
// are coalesced and the 4-byte stores of consecutive w fall into the same sectors of L2.
static __global__ void __launch_bounds__(128) moves_gather_kernel(const uint32_t *mvg, uint32_t stride,
                                                                  const uint32_t *aln_len, const uint64_t *mvoff,
                                                                  uint32_t first, uint32_t count, uint32_t *out) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t nw = 0;
  uint64_t o = 0;
  if (slot < count) {
    nw = (aln_len[first + slot] + 14u) >> 4;
    o = mvoff[first + slot];
  }
  uint32_t wmax = nw;
#pragma unroll
  for (int k = 16; k > 0; k >>= 1) wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, k));
  for (uint32_t w = 0; w < wmax; ++w)
    if (w < nw) out[o + w] = mvg[(size_t)w * stride + slot];
}

// Move words from aligned strings (every traceback that is not the move-string one): one warp per
// pair, a lane builds one word from 16 columns read backwards from the column before the seed.
static __global__ void __launch_bounds__(128) strings_to_moves_kernel(const uint8_t *q_aln, const uint8_t *t_aln,
                                                                      const uint64_t *aln_off, const uint32_t *aln_len,
                                                                      const uint64_t *mvoff, uint64_t count,
                                                                      uint32_t *out) {
  const uint64_t pair = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (pair >= count) return;
  const uint32_t L = aln_len[pair];
  if (L < 2) return;
  const uint32_t K = L - 1, nw = (K + 15) >> 4;
  const uint64_t last = aln_off[pair] + L - 2;  // column of move 0
  for (uint32_t w = lane; w < nw; w += 32) {
    uint32_t word = 0;
    for (uint32_t j = 0; j < 16 && w * 16 + j < K; ++j) {
      const uint64_t pos = last - (w * 16 + j);
      const uint32_t code = q_aln[pos] == kBlank ? DIR_TOP : (t_aln[pos] == kBlank ? DIR_LEFT : DIR_DIAGONAL);
      word |= code << (2 * j);
    }
    out[mvoff[pair] + w] = word;
  }
}

// ---- compact input format (alnr_batch.q2 / .t2) ---------------------------------------------------
// 2-bit codes -> byte codes for the sequences [first, first + count).  woff == nullptr: every
// sequence has nw words (uniform lengths) and one thread expands one word; else one warp per sequence.
static __device__ __forceinline__ void unpack_word(uint32_t word, uint8_t *dst, uint64_t left) {
  if (left >= 16 && ((uintptr_t)dst & 1) == 0) {
    uint16_t *d2 = reinterpret_cast<uint16_t *>(dst);
#pragma unroll
    for (int j = 0; j < 8; ++j) d2[j] = (uint16_t)(((word >> (4 * j)) & 3u) | (((word >> (4 * j + 2)) & 3u) << 8));
  } else {
    for (uint32_t j = 0; j < 16 && j < left; ++j) dst[j] = (uint8_t)((word >> (2 * j)) & 3u);
  }
}

static __global__ void __launch_bounds__(256) unpack2bit_kernel(const uint32_t *in, const uint64_t *woff, uint32_t nw,
                                                                const uint64_t *off, uint64_t first, uint64_t count,
                                                                uint8_t *out) {
  const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (!woff) {
    const uint64_t s = gid / nw;
    const uint32_t k = (uint32_t)(gid % nw);
    if (s >= count) return;
    const uint64_t b = off[first + s], e = off[first + s + 1];
    if ((uint64_t)k * 16 < e - b) unpack_word(in[(first + s) * nw + k], out + b + (uint64_t)k * 16, e - b - (uint64_t)k * 16);
  } else {
    const uint64_t s = gid >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (s >= count) return;
    const uint64_t b = off[first + s], e = off[first + s + 1], w0 = woff[first + s];
    for (uint64_t k = lane; k * 16 < e - b; k += 32) unpack_word(in[w0 + k], out + b + k * 16, e - b - k * 16);
  }
}

}  // namespace alnr

namespace alnr {

// Traceback of ONE long pair through the "blocked" D layout (band pipeline kernel, 2.5 GB of
// directions for 10^5 x 10^5 cells): a single dependent chain of ~n+m loads that mostly miss L2.
// One warp walks the path in lockstep (every lane holds the same (cy, cx)); before each stretch
// the 32 lanes fetch the words the path is expected to cross in the next 16 strips — two words
// per strip around the diagonal through the current cell — in one batch of independent loads,
// and the walk then takes its words from the owning lane by shuffle.  A word that was not
// predicted (the path drifted by indels) is loaded directly.  Output as traceback_kernel.
// Checkpointed mode (ck != nullptr; one pair): the direction scratch only holds the strips
// [strip0, strip0 + dl.nstrips) of one GROUP of bands, re-filled from the kept band boundary rows
// (wave.cu).  The walk resumes from *ck, stops when the path climbs above row y_stop (the rows of
// the groups still to be re-filled) and leaves its state in *ck for the next group's launch.
struct CkptState {
  uint32_t cy, cx;          // current cell of the walk
  unsigned long long pos;   // next output byte (the strings grow downwards from the slot's end)
  uint32_t done, started;
};

static __global__ void ckpt_init_kernel(TraceArgs a, CkptState *ck) {
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first] : a.cv.first;
  const uint32_t sp = a.cv.seq(pair);
  const uint64_t qo = a.cv.q_off[sp], to = a.cv.t_off[sp];
  ck->cy = a.fo.start[2 * (size_t)pair];
  ck->cx = a.fo.start[2 * (size_t)pair + 1];
  ck->pos = qo + to + pair + (a.cv.q_off[sp + 1] - qo) + (a.cv.t_off[sp + 1] - to) + 1;
  ck->done = 0;  // a rejected pair is finished by the first walk launch
  ck->started = 0;
}

static __global__ void __launch_bounds__(32) traceback_warp_kernel(TraceArgs a, CkptState *ck = nullptr,
                                                                   uint32_t y_stop = 0, uint32_t strip0 = 0) {
  const uint32_t slot = blockIdx.x, lane = threadIdx.x;
  if (slot >= a.cv.count) return;
  if (ck && ck->done) return;
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first + slot] : a.cv.first + slot;
  const uint32_t sp = a.cv.seq(pair);
  const uint64_t qo = a.cv.q_off[sp], to = a.cv.t_off[sp];
  const uint32_t n = (uint32_t)(a.cv.q_off[sp + 1] - qo);
  const uint32_t m = (uint32_t)(a.cv.t_off[sp + 1] - to);
  const uint64_t slot_end = qo + to + pair + n + m + 1;
  if (a.fo.status[pair] != 0) {
    if (lane == 0) {
      if (a.to.aln_len) a.to.aln_len[pair] = 0;
      if (a.to.aln_off) a.to.aln_off[pair] = slot_end;
      if (a.to.coords) for (int k = 0; k < 4; ++k) a.to.coords[4 * (size_t)pair + k] = 0;
      if (ck) ck->done = 1;
    }
    return;
  }
  const uint8_t *q = a.cv.q + qo, *t = a.cv.t + to;
  const bool glob = a.mode == 0;
  // R and C are powers of two (R * C = 16): shifts and masks instead of divisions on the chain
  const uint32_t R = a.dl.rows_per_word, C = a.dl.blocked;
  const uint32_t rsh = __ffs(R) - 1, csh = __ffs(C) - 1, rmask = R - 1, cmask = C - 1;
  const uint32_t sy = a.fo.start[2 * (size_t)pair], sx = a.fo.start[2 * (size_t)pair + 1];
  uint32_t cy = sy, cx = sx;
  uint64_t pos = slot_end;
  bool resumed = false;
  if (ck) {
    cy = ck->cy; cx = ck->cx; pos = ck->pos; resumed = ck->started != 0;
    // the path is not inside this group's rows (local mode: the arg-max sits higher up): nothing to do
    if (cy && cx && ((cy - 1) >> rsh) >= strip0 + a.dl.nstrips) return;
  }
  const bool wr = lane == 0 && a.to.q_aln != nullptr && a.to.t_aln != nullptr;
  auto emit = [&](uint8_t qc, uint8_t tc) {
    --pos;
    if (wr) { a.to.q_aln[pos] = qc; a.to.t_aln[pos] = tc; }
  };
  if (!resumed) emit(q[cx - 1], t[cy - 1]);  // seed (simple/mod.rs:102-105, :213-216)
  bool done = false, paused = false;
  while (!done && !paused) {
    // ---- batch of predicted words: strip base - k, k = 0..15, two words each ---------------
    uint32_t my_val = 0;
    const uint32_t base = cy ? (cy - 1) >> rsh : 0;
    const uint32_t rows_here = cy ? ((cy - 1) & rmask) + 1 : 0;  // rows left in the current strip
    const int64_t cx0 = cx;                                // column at the start of this batch
    auto pred_hi = [&](uint32_t k) -> int64_t {           // word column where the path enters strip base - k
      const int64_t px = cx0 - 1 - (k ? (int64_t)rows_here + (int64_t)(k - 1) * R : 0);
      return px >= 0 ? (px >> csh) : -1;
    };
    {
      const uint32_t k = lane >> 1;
      const int64_t wc = pred_hi(k) - (lane & 1);
      if (cy && cx && k <= base && base - k >= strip0 && wc >= 0)
        my_val = __ldcs(a.D + ((uint64_t)slot * a.dl.nstrips + (base - k - strip0)) * a.dl.ncols + (uint64_t)wc);
    }
    // ---- walk until the path leaves the 16 predicted strips ---------------------------------
    uint32_t cur_strip = 0xffffffffu, cur_wc = 0xffffffffu, cur_w = 0;
    for (;;) {
      int d;
      if (cy == 0 || cx == 0) {
        if (!glob || (cy == 0 && cx == 0)) d = DIR_BEGINNING;
        else d = (cy == 0) ? DIR_LEFT : DIR_TOP;  // simple/mod.rs:59-67
      } else {
        if (ck && cy <= y_stop) { paused = true; break; }  // the rows above belong to the next group
        const uint32_t strip = (cy - 1) >> rsh, wc = (cx - 1) >> csh;
        if (strip != cur_strip || wc != cur_wc) {
          const uint32_t k = base - strip;
          if (k >= 16) break;  // next batch
          const int64_t hi = pred_hi(k);
          const uint32_t src = 2 * k + (((int64_t)wc == hi) ? 0u : 1u);
          const uint32_t v = __shfl_sync(0xffffffffu, my_val, src);
          if ((int64_t)wc == hi || (int64_t)wc == hi - 1) cur_w = v;
          else cur_w = a.D[((uint64_t)slot * a.dl.nstrips + (strip - strip0)) * a.dl.ncols + wc];  // not predicted
          cur_strip = strip; cur_wc = wc;
        }
        d = (cur_w >> (2 * (((cx - 1) & cmask) << rsh) + 2 * ((cy - 1) & rmask))) & 3u;
      }
      if (d == DIR_BEGINNING) { done = true; break; }
      if (d == DIR_TOP) { emit(kBlank, t[cy - 1]); --cy; }
      else if (d == DIR_LEFT) { emit(q[cx - 1], kBlank); --cx; }
      else { emit(q[cx - 1], t[cy - 1]); --cx; --cy; }
    }
  }
  if (lane == 0 && ck) { ck->cy = cy; ck->cx = cx; ck->pos = pos; ck->started = 1; ck->done = done ? 1u : 0u; }
  if (lane == 0 && (done || !ck)) {
    if (a.to.aln_len) a.to.aln_len[pair] = (uint32_t)(slot_end - pos);
    if (a.to.aln_off) a.to.aln_off[pair] = pos;
    if (a.to.coords) {
      uint64_t *c = a.to.coords + 4 * (size_t)pair;
      if (glob) { c[0] = 1; c[1] = n; c[2] = 1; c[3] = m; }                      // simple/mod.rs:137
      else { c[0] = cx + 1; c[1] = sx + 1; c[2] = cy + 1; c[3] = sy + 1; }        // :254-257
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Two-phase traceback of ONE long pair through the "blocked" D layout with 4 x 4 tiles per word
// (tile pipeline kernel).  The walk is one dependent chain of n + m steps; everything that is not
// on that chain is moved off it:
//   phase W (trace_walk_coop_kernel, one warp): the walk happens inside a WINDOW of 32 strips x 64
//     words (128 rows x 256 columns, 8 KB) held in shared memory and emits only 2-bit moves (16 per
//     word); no sequence loads, no alignment stores.  All 32 lanes fetch a window with 16-byte
//     cp.async; while the window is walked, the one the path is predicted to enter next (straight
//     up the diagonal, 16 words of slack to the right, 47 to the left) is already in flight into the
//     second buffer.  A path that leaves a window elsewhere (long gap runs) gets a window loaded at
//     its cell.  (The first version walked the window with lane 0 alone: 8.9 ms per 10^5 x 10^5
//     against 1.7 ms for the cooperative walk below; removed in round 2.)
//   phase E (trace_expand_long_kernel, one thread per move word): re-walks its 16 moves from the
//     recorded cell and writes the two aligned strings, right-aligned in the pair's slot exactly
//     as traceback_kernel does (seed pair first: simple/mod.rs:102-105, :213-216).
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kWinStrips = 32, kWinWords = 64;

struct LongTrace {
  uint32_t *moves;  // [nmw] 2-bit moves, move i at bits 2 * (i % 16) of word i / 16
  uint32_t *cell;   // [2 * nmw] (cy, cx) before the first move of every word
  uint32_t *tail;   // [0] number of moves, [1] cy and [2] cx where the walk ended
};

__device__ __forceinline__ void win_cp16(uint32_t smem_addr, const void *g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(g) : "memory");
}

// window with bottom strip s_bot and rightmost word w_hi: strips s_bot - 31 .. s_bot (row r of the
// buffer = strip s_bot - r), words w_lo .. w_lo + 63 with w_lo a multiple of 4 (ncols is one too)
__device__ __forceinline__ uint32_t win_lo(uint32_t w_hi) {
  const int lo = ((int)w_hi - (int)kWinWords + 1 + 3) & ~3;
  return lo > 0 ? (uint32_t)lo : 0u;
}

__device__ __forceinline__ void win_fetch(uint32_t *buf, const uint32_t *D, const DLayout &dl, uint32_t s_bot,
                                          uint32_t w_lo, uint32_t lane) {
  const uint32_t piece = lane & 15, half = lane >> 4;
  const uint32_t w = w_lo + 4 * piece;
#pragma unroll 4
  for (uint32_t r2 = 0; r2 < kWinStrips / 2; ++r2) {
    const uint32_t r = 2 * r2 + half;
    if (r <= s_bot && w < dl.ncols)
      win_cp16((uint32_t)__cvta_generic_to_shared(buf + r * kWinWords + 4 * piece),
               D + (uint64_t)(s_bot - r) * dl.ncols + w);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}


// Cooperative window walk: the 32 strips of a window are walked by the 32 lanes AT ONCE.  Lane l takes strip s_bot - l
// and starts on the bottom row of its strip at a GUESSED column — where the path would arrive if
// it kept climbing the diagonal from the last known cell.  The guesses are then stitched: lane l's
// walk is the real path iff lane l-1's walk is and leaves its strip at the column lane l started
// from.  The first lane whose guess was wrong gets the true column, the lanes above it are
// re-guessed from there, and only those walk again; an indel costs one more round instead of
// making the walk sequential.  (A round confirms at least one more strip, so the worst case is
// the sequential walk.)  A strip walk stops after 32 moves, at Beginning, or at the window's left
// edge; the chain then ends at that cell and the outer loop fetches a window there.
// Moves only: the start cells of the move words come from trace_cells_kernel.
constexpr uint32_t kStripMoves = 32;  // longest walk of a strip entered at a known cell
constexpr uint32_t kGuessMoves = 8;   // longest walk from a guessed cell (it is repeated in full once the cell is confirmed)

static __global__ void __launch_bounds__(32) trace_walk_coop_kernel(TraceArgs a, LongTrace lt) {
  __shared__ __align__(16) uint32_t win[2][kWinStrips * kWinWords];
  __shared__ uint32_t stage[72];  // move bits being appended; stage[0] starts as the open move word
  const uint32_t lane = threadIdx.x;
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first] : a.cv.first;
  if (a.fo.status[pair] != 0) {
    if (lane == 0) { lt.tail[0] = 0; lt.tail[1] = 0; lt.tail[2] = 0; }
    return;
  }
  const bool glob = a.mode == 0;
  uint32_t cy = a.fo.start[2 * (size_t)pair], cx = a.fo.start[2 * (size_t)pair + 1];
  uint32_t cnt = 0;
  int cur = 0;
  for (uint32_t i = lane; i < 72; i += 32) stage[i] = 0;
  __syncwarp();
  uint32_t s_bot = 0, w_lo = 0, seq_left = 0;
  uint32_t ns_bot = 0, nw_lo = 0, nw_hi = 0;  // the prefetched window (pf_have)
  bool have = false, done = false, pf_have = false;
  while (!done) {
    const bool inside = cy && cx;
    if (!inside) {  // border: Beginning in local mode, a run to (0, 0) in global mode (simple/mod.rs:59-67)
      if (lane == 0 && glob) {
        uint32_t acc = stage[0];
        auto put = [&](uint32_t d) {
          acc |= d << (2 * (cnt & 15));
          ++cnt;
          if ((cnt & 15) == 0) { lt.moves[(cnt >> 4) - 1] = acc; acc = 0; }
        };
        while (cx) { --cx; put(DIR_LEFT); }
        while (cy) { --cy; put(DIR_TOP); }
        stage[0] = acc;
      }
      cnt = __shfl_sync(0xffffffffu, cnt, 0);
      cy = __shfl_sync(0xffffffffu, cy, 0);
      cx = __shfl_sync(0xffffffffu, cx, 0);
      __syncwarp();
      break;
    }
    const uint32_t strip = (cy - 1) >> 2, wc = (cx - 1) >> 2;
    if (!have) {  // no usable window: fetch one at the current cell
      s_bot = strip; w_lo = win_lo(wc);
      win_fetch(win[cur], a.D, a.dl, s_bot, w_lo, lane);
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncwarp();
      have = true;
    }
    const uint32_t s_top = s_bot >= kWinStrips - 1 ? s_bot - (kWinStrips - 1) : 0;
    // predicted next window: the path climbs the rest of this window along the diagonal
    // (kept across iterations that stay in the same window)
    if (!pf_have && s_top > 0) {
      const uint32_t climb = strip - s_top + 1;
      if (wc > climb) {
        ns_bot = s_top - 1;
        nw_hi = min(wc - climb + 16, a.dl.ncols - 1);
        nw_lo = win_lo(nw_hi);
        win_fetch(win[cur ^ 1], a.D, a.dl, ns_bot, nw_lo, lane);
        pf_have = true;
      }
    }
    // ---- the window's strips, one per lane -------------------------------------------------
    const uint32_t *wb = win[cur];
    const uint32_t lmax = strip - s_top;          // lanes 0 .. lmax have a strip (strip, not s_bot: the
    const bool mine = lane <= lmax;               //  path may enter a prefetched window below its top)
    const uint32_t my_strip = strip - (mine ? lane : 0);
    const uint32_t ybase = my_strip * 4;          // the walk of this lane ends when y == ybase
    uint32_t ey = lane == 0 ? cy : ybase + 4;     // entry cell of the lane
    uint32_t ex = lane == 0 ? cx : (cx > cy - ey ? cx - (cy - ey) : 0u);
    uint32_t confirmed = 0;                       // lanes <= confirmed start from the true cell
    bool redo = true;                             // this lane has to walk (again)
    unsigned long long mv = 0;
    // st: 0 left the strip at the top, 1 Beginning, 2 window edge / 32 moves, 3 a guessed walk cut short
    uint32_t k = 0, oy = 0, ox = 0, st = 0;
    uint32_t c_end = 0, rounds = 0;
    if (seq_left) {  // speculation kept failing (unrelated sequences): lane 0 walks this window alone
      --seq_left;
      if (lane == 0) {
        uint32_t acc = stage[0];
        uint32_t y = cy, x = cx;
        st = 0;
        for (;;) {
          if (y == 0 || x == 0) break;  // the border run is the outer loop's
          const uint32_t sy = (y - 1) >> 2, ww = (x - 1) >> 2;
          if (sy < s_top || ww < w_lo) break;
          const uint32_t w = wb[(s_bot - sy) * kWinWords + (ww - w_lo)];
          int ly = (int)((y - 1) & 3), lx = (int)((x - 1) & 3);
          bool begin = false;
          do {
            const uint32_t d = (w >> (8 * lx + 2 * ly)) & 3u;
            if (d == DIR_BEGINNING) { begin = true; break; }
            ly -= (d != DIR_LEFT);
            lx -= (d != DIR_TOP);
            acc |= d << (2 * (cnt & 15));
            ++cnt;
            if ((cnt & 15) == 0) { lt.moves[(cnt >> 4) - 1] = acc; acc = 0; }
          } while (ly >= 0 && lx >= 0);
          y = sy * 4 + 1 + ly; x = ww * 4 + 1 + lx;
          if (begin) { st = 1; break; }
        }
        stage[0] = acc;
        oy = y; ox = x;
      }
      cnt = __shfl_sync(0xffffffffu, cnt, 0);
      cy = __shfl_sync(0xffffffffu, oy, 0);
      cx = __shfl_sync(0xffffffffu, ox, 0);
      done = __shfl_sync(0xffffffffu, st, 0) == 1 || (!glob && (cy == 0 || cx == 0));
    } else {
    for (;;) {
      ++rounds;
      if (mine && redo) {
        uint32_t y = ey, x = ex;
        mv = 0; k = 0;
        const uint32_t cap = lane <= confirmed ? kStripMoves : kGuessMoves;
        for (;;) {
          uint32_t d;
          if (x == 0) {
            if (!glob) { st = 1; break; }
            d = DIR_TOP;
          } else {
            const uint32_t ww = (x - 1) >> 2;
            if (ww < w_lo) { st = 2; break; }
            const uint32_t w = wb[(s_bot - my_strip) * kWinWords + (ww - w_lo)];
            d = (w >> (8 * ((x - 1) & 3) + 2 * ((y - 1) & 3))) & 3u;
            if (d == DIR_BEGINNING) { st = 1; break; }
          }
          if (k == cap) { st = cap == kStripMoves ? 2 : 3; break; }
          mv |= (unsigned long long)d << (2 * k);
          ++k;
          y -= (d != DIR_LEFT);
          x -= (d != DIR_TOP);
          if (y == ybase) { st = 0; break; }
        }
        oy = y; ox = x;
      }
      __syncwarp();
      // stitch: lane l is on the path iff lane l-1 is, left its strip at the top, and did so at lane l's entry column
      const uint32_t pox = __shfl_up_sync(0xffffffffu, ox, 1), pst = __shfl_up_sync(0xffffffffu, st, 1);
      const bool bad = mine && lane > confirmed && (pst != 0 || pox != ex || st == 3);
      const uint32_t badmask = __ballot_sync(0xffffffffu, bad);
      const uint32_t first_bad = badmask ? (uint32_t)__ffs(badmask) - 1 : lmax + 1;
      confirmed = first_bad - 1;
      const uint32_t cst = __shfl_sync(0xffffffffu, st, confirmed);
      if (first_bad > lmax || cst != 0) {
        c_end = confirmed;
        if (cst == 2 && confirmed < 2) seq_left = 32;  // a gap run longer than a strip walk: no diagonal to guess along
        break;
      }
      if (rounds > 8) {  // one strip per round: the guesses are not paying (unrelated sequences) — lane 0
        c_end = confirmed; seq_left = 32; break;  // walks the rest of this window and the next 31 alone
      }
      // lane first_bad starts from the true column, the lanes above it from the diagonal through it
      const uint32_t tx = __shfl_sync(0xffffffffu, ox, confirmed), ty = __shfl_sync(0xffffffffu, oy, confirmed);
      redo = mine && lane >= first_bad;
      if (redo) ex = tx > ty - ey ? tx - (ty - ey) : 0u;
      confirmed = first_bad;
    }
    // ---- append the moves of lanes 0 .. c_end ----------------------------------------------
    {
      const uint32_t kk = (mine && lane <= c_end) ? k : 0u;
      uint32_t incl = kk;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)lane >= o) incl += v;
      }
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      if (kk) {
        const uint32_t bitpos = 2 * ((cnt & 15) + incl - kk);
        const uint32_t wi = bitpos >> 5, sh = bitpos & 31;
        const uint32_t lo = (uint32_t)mv, hi = (uint32_t)(mv >> 32);
        atomicOr(&stage[wi], lo << sh);
        atomicOr(&stage[wi + 1], sh ? ((lo >> (32 - sh)) | (hi << sh)) : hi);
        if (sh) atomicOr(&stage[wi + 2], hi >> (32 - sh));
      }
      __syncwarp();
      const uint32_t full = ((cnt & 15) + total) >> 4;
      for (uint32_t i = lane; i < full; i += 32) lt.moves[(cnt >> 4) + i] = stage[i];
      const uint32_t open = stage[full];
      __syncwarp();
      for (uint32_t i = lane; i < 72; i += 32) stage[i] = 0;
      __syncwarp();
      if (lane == 0) stage[0] = open;
      __syncwarp();
      cnt += total;
    }
    cy = __shfl_sync(0xffffffffu, oy, c_end);
    cx = __shfl_sync(0xffffffffu, ox, c_end);
    done = __shfl_sync(0xffffffffu, st, c_end) == 1;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    have = false;
    if (!done && cy && cx) {
      const uint32_t s2 = (cy - 1) >> 2, w2 = (cx - 1) >> 2;
      const uint32_t ns_top = ns_bot >= kWinStrips - 1 ? ns_bot - (kWinStrips - 1) : 0;
      if (s2 >= s_top && s2 <= s_bot && w2 >= w_lo) {
        have = true;  // a walk cut short (32 moves, speculation given up) continues in the same window
      } else if (pf_have && s2 <= ns_bot && s2 >= ns_top && w2 >= nw_lo && w2 <= nw_hi) {
        cur ^= 1; s_bot = ns_bot; w_lo = nw_lo; have = true; pf_have = false;
      }
    }
    if (!have) pf_have = false;
  }
  if (lane == 0) {
    if (cnt & 15) lt.moves[cnt >> 4] = stage[0];
    lt.tail[0] = cnt; lt.tail[1] = cy; lt.tail[2] = cx;
  }
}

// Start cell of every 16-move word by a prefix sum over the move string (one block): a word with
// L Left and G Diagonal moves among its c moves climbs c - L rows and G + L columns.
static __global__ void __launch_bounds__(1024) trace_cells_kernel(TraceArgs a, LongTrace lt) {
  __shared__ uint32_t sy[1024], sx[1024];
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first] : a.cv.first;
  if (a.fo.status[pair] != 0) return;
  const uint32_t cnt = lt.tail[0], nw = (cnt + 15) >> 4;
  const uint32_t per = (nw + 1023) / 1024, w0 = threadIdx.x * per, w1 = min(nw, w0 + per);
  auto climb = [&](uint32_t i, uint32_t &dy, uint32_t &dx) {
    const uint32_t w = lt.moves[i], c = min(16u, cnt - 16 * i);
    const uint32_t L = __popc(w & 0x55555555u), G = __popc((w >> 1) & 0x55555555u);
    dy = c - L; dx = G + L;
  };
  uint32_t ty = 0, tx = 0;
  for (uint32_t i = w0; i < w1; ++i) { uint32_t dy, dx; climb(i, dy, dx); ty += dy; tx += dx; }
  sy[threadIdx.x] = ty; sx[threadIdx.x] = tx;
  __syncthreads();
  for (uint32_t o = 1; o < 1024; o <<= 1) {
    const uint32_t vy = threadIdx.x >= o ? sy[threadIdx.x - o] : 0, vx = threadIdx.x >= o ? sx[threadIdx.x - o] : 0;
    __syncthreads();
    sy[threadIdx.x] += vy; sx[threadIdx.x] += vx;
    __syncthreads();
  }
  uint32_t cy = a.fo.start[2 * (size_t)pair] - (sy[threadIdx.x] - ty);
  uint32_t cx = a.fo.start[2 * (size_t)pair + 1] - (sx[threadIdx.x] - tx);
  for (uint32_t i = w0; i < w1; ++i) {
    lt.cell[2 * i] = cy; lt.cell[2 * i + 1] = cx;
    uint32_t dy, dx; climb(i, dy, dx);
    cy -= dy; cx -= dx;
  }
}

static __global__ void __launch_bounds__(256) trace_expand_long_kernel(TraceArgs a, LongTrace lt) {
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first] : a.cv.first;
  const uint32_t sp = a.cv.seq(pair);
  const uint64_t qo = a.cv.q_off[sp], to = a.cv.t_off[sp];
  const uint32_t n = (uint32_t)(a.cv.q_off[sp + 1] - qo);
  const uint32_t m = (uint32_t)(a.cv.t_off[sp + 1] - to);
  const uint64_t slot_end = qo + to + pair + n + m + 1;
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (a.fo.status[pair] != 0) {
    if (gid == 0) {
      if (a.to.aln_len) a.to.aln_len[pair] = 0;
      if (a.to.aln_off) a.to.aln_off[pair] = slot_end;
      if (a.to.coords) for (int k = 0; k < 4; ++k) a.to.coords[4 * (size_t)pair + k] = 0;
    }
    return;
  }
  const uint8_t *q = a.cv.q + qo, *t = a.cv.t + to;
  const uint32_t cnt = lt.tail[0];
  const uint32_t sy = a.fo.start[2 * (size_t)pair], sx = a.fo.start[2 * (size_t)pair + 1];
  const bool wr = a.to.q_aln != nullptr && a.to.t_aln != nullptr;
  if (gid == 0) {
    if (wr) { a.to.q_aln[slot_end - 1] = q[sx - 1]; a.to.t_aln[slot_end - 1] = t[sy - 1]; }  // seed pair
    if (a.to.aln_len) a.to.aln_len[pair] = cnt + 1;
    if (a.to.aln_off) a.to.aln_off[pair] = slot_end - 1 - cnt;
    if (a.to.coords) {
      uint64_t *c = a.to.coords + 4 * (size_t)pair;
      if (a.mode == 0) { c[0] = 1; c[1] = n; c[2] = 1; c[3] = m; }                          // simple/mod.rs:137
      else { c[0] = lt.tail[2] + 1; c[1] = sx + 1; c[2] = lt.tail[1] + 1; c[3] = sy + 1; }  // :254-257
    }
  }
  const uint32_t first = gid * 16;
  if (first >= cnt || !wr) return;
  uint32_t w = lt.moves[gid];
  uint32_t cy = lt.cell[2 * gid], cx = lt.cell[2 * gid + 1];
  const uint32_t k_end = min(16u, cnt - first);
  uint64_t pos = slot_end - 2 - first;  // move i lands at slot_end - 2 - i
  for (uint32_t k = 0; k < k_end; ++k, --pos, w >>= 2) {
    const uint32_t d = w & 3u;
    a.to.q_aln[pos] = d == DIR_TOP ? (uint8_t)kBlank : q[cx - 1];
    a.to.t_aln[pos] = d == DIR_LEFT ? (uint8_t)kBlank : t[cy - 1];
    cy -= (d != DIR_LEFT);
    cx -= (d != DIR_TOP);
  }
}

// Traceback of the matrix search (heur.cu): one WARP per item through the "skewed" D layout of the
// warp fill kernel; only the coords and the frequency matrix of the alignment (alignment.rs:13-23)
// are produced.  Every lane holds the same (cy, cx); before each stretch the 32 lanes fetch a
// window of 2 strips x 16 columns up-left of the current cell in one batch of independent loads
// and the walk takes its words from the owning lane by shuffle, so the chain pays one L2 latency
// per ~14 steps instead of one per step.  The counts are kept in shared memory (lane 0 adds) and
// written once, coalesced; the two sequences are staged in shared memory too.
// grid: ceil(count / 4) blocks of 128 threads; smem: 4 * (V*V*8 + pad8(n_max) + pad8(m_max)).
static __global__ void __launch_bounds__(128) traceback_freq_warp_kernel(TraceArgs a, uint32_t n_max, uint32_t m_max) {
  extern __shared__ double trace_smem[];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t slot = blockIdx.x * 4 + warp;
  if (slot >= a.cv.count) return;
  const uint32_t pair = a.cv.order ? a.cv.order[a.cv.first + slot] : a.cv.first + slot;
  const int VV = a.V * a.V;
  const uint32_t nq = (n_max + 7) / 8 * 8, nt = (m_max + 7) / 8 * 8;
  double *sf = trace_smem + (size_t)warp * (VV + (nq + nt) / 8);
  uint8_t *sq = reinterpret_cast<uint8_t *>(sf + VV), *st = sq + nq;
  double *freq = a.to.freq + (size_t)pair * VV;
  if (a.fo.status[pair] != 0) {
    for (int i = lane; i < VV; i += 32) freq[i] = 0.0;
    if (lane == 0 && a.to.coords) for (int k = 0; k < 4; ++k) a.to.coords[4 * (size_t)pair + k] = 0;
    return;
  }
  const uint32_t sp = a.cv.seq(pair);
  const uint8_t *q, *t;
  uint32_t n, m;
  a.cv.pair_dims(sp, q, t, n, m);
  for (int i = lane; i < VV; i += 32) sf[i] = 0.0;
  for (uint32_t i = lane; i < n; i += 32) sq[i] = q[i];
  for (uint32_t i = lane; i < m; i += 32) st[i] = t[i];
  __syncwarp();
  const uint32_t R = a.dl.rows_per_word;
  const uint32_t sy = a.fo.start[2 * (size_t)pair], sx = a.fo.start[2 * (size_t)pair + 1];
  uint32_t cy = sy, cx = sx;
  const uint32_t *Dslot = a.D + (uint64_t)slot * a.dl.nstrips * (a.dl.ncols + 31);
  auto word_at = [&](uint32_t strip, uint32_t xm1) {
    const uint32_t band = strip >> 5, l = strip & 31;
    return Dslot[((uint64_t)band * (a.dl.ncols + 31) + xm1 + l) * 32ull + l];
  };
  auto emit = [&](uint32_t qc, uint32_t tc) {
    if (lane == 0 && qc != (uint32_t)kBlank && tc != (uint32_t)kBlank) sf[tc * a.V + qc] += 1.0;  // alignment.rs:16-20
  };
  emit(sq[cx - 1], st[cy - 1]);  // seed (simple/mod.rs:213-216)
  uint32_t strip = (cy - 1) / R, r = (cy - 1) % R;  // kept incrementally: no division on the chain
  bool done = false;
  while (!done) {
    // ---- window: strips strip0, strip0 - 1; columns cx0 - 1 ... cx0 - 16 (0-based xm1) -------
    const uint32_t strip0 = strip, x0 = cx;  // x0 >= 1 here
    uint32_t my_w = 0;
    {
      const uint32_t j = lane >> 4, k = lane & 15;
      if (j <= strip0 && k < x0) my_w = word_at(strip0 - j, x0 - 1 - k);
    }
    for (;;) {
      if (cy == 0 || cx == 0) { done = true; break; }  // local mode: row 0 / column 0 are Beginning
      const uint32_t j = strip0 - strip, k = x0 - cx;
      if (j >= 2 || k >= 16) break;  // next window
      const uint32_t w = __shfl_sync(0xffffffffu, my_w, j * 16 + k);
      const uint32_t d = (w >> (2 * r)) & 3u;
      if (d == DIR_BEGINNING) { done = true; break; }
      const bool up = d != DIR_LEFT, left = d != DIR_TOP;
      emit(left ? sq[cx - 1] : (uint32_t)kBlank, up ? st[cy - 1] : (uint32_t)kBlank);
      if (up) { --cy; if (r == 0) { r = R - 1; --strip; } else --r; }
      if (left) --cx;
    }
  }
  __syncwarp();
  for (int i = lane; i < VV; i += 32) freq[i] = sf[i];
  if (lane == 0 && a.to.coords) {
    uint64_t *c = a.to.coords + 4 * (size_t)pair;
    c[0] = cx + 1; c[1] = sx + 1; c[2] = cy + 1; c[3] = sy + 1;  // simple/mod.rs:254-257
  }
}

}  // namespace alnr
