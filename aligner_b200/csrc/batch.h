// batch.h — device-resident batch of independent pairs (host-side class behind
// alnr_batch_* and alnr_align_batch / alnr_align_pair).
#pragma once
#include <vector>

#include "common.cuh"
#include "host_util.h"

namespace alnr {

struct Chunk {
  uint32_t first, count;      // slots
  uint32_t m_max, n_max;      // chunk-wide maxima
  uint64_t q0, q1, t0, t1;    // symbol ranges of the chunk (identity order only)
  bool wave = false;          // one long pair: intra-pair wavefront kernel
  uint32_t wave_pair = 0;     // its pair index; q0/t0 = its symbol offsets
  uint32_t ckpt_group = 0;    // > 0: checkpointed traceback, the D scratch holds this many bands
};

struct alnr_batch_impl {
  // configuration
  int mode = 0, V = 0, device = 0, numeric = 0;
  double del = 0, ext = 0;
  std::vector<double> matrix;
  bool want_matrices = false;   // single-pair calls: also produce H and D (reference layout)
  // score-only calls (alnr_score_batch; statistics/mod.rs:255-292 keeps only alignment.f): no D
  // scratch, no traceback, no alignment buffers
  bool score_only = false;
  // compact result format (alnr_batch_out.moves): the traceback stops at the 2-bit move strings
  bool moves_out = false;
  // every chunk takes the move-string traceback (uniform short pairs): moves are gathered chunk by
  // chunk; otherwise one pass over the whole batch derives them from the aligned strings
  bool moves_direct = false;
  bool want_coords = true;      // false: the caller passed no coords array (host -> host call)
  // compact input format (alnr_batch.q2 / .t2): 2-bit DNA, 16 codes per word, pair-major
  bool packed_in = false;
  uint64_t q2_words = 0, t2_words = 0, mv_capacity = 0;
  std::vector<uint64_t> h_qwoff, h_twoff;  // ragged packed input: word offset of every sequence

  // shape
  uint64_t count = 0, q_total = 0, t_total = 0;
  uint32_t n_max = 0, m_max = 0;
  bool uniform = true;
  std::vector<Chunk> chunks;

  // device state
  cudaStream_t stream = nullptr;    // phase 1 of every chunk: validation, packing, main fill pass
  // Phase 2 of a chunk (speculation re-runs, exact leftovers, traceback) runs on a second stream
  // so that its thin launches overlap the next chunk's main pass; the D / row / list scratch is
  // double-buffered by chunk parity.  `stream` waits for the last phase 2 at the end of run(), so
  // synchronising `stream` still means "everything is done".
  cudaStream_t stream2 = nullptr;
  cudaStream_t s2 = nullptr;        // stream phase 2 is enqueued on right now (stream2, or stream when serial)
  std::vector<cudaEvent_t> ev_p1, ev_p2;  // per chunk: phase 1 done, phase 2 done
  int par = 0;                      // parity of the chunk being enqueued
  uint64_t d_words = 0, hcol_elems = 0, rerun_words = 0;  // size of ONE parity copy
  uint32_t *Dp() const { return score_only ? nullptr : d_D.p + (size_t)par * d_words; }
  int *hcol_i() const { return d_hcol_i.p ? d_hcol_i.p + (size_t)par * hcol_elems : nullptr; }
  double *hcol_d() const { return d_hcol_d.p ? d_hcol_d.p + (size_t)par * hcol_elems : nullptr; }
  int phase1(size_t ci);            // on `stream`
  int phase2(size_t ci);            // on `s2`
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
  DevBuf<uint8_t> d_q, d_t, d_qaln, d_taln, d_Dout;
  DevBuf<uint64_t> d_qoff, d_toff, d_coords, d_alnoff;
  DevBuf<uint64_t> d_desc;      // compact traceback: [ticket][one scan descriptor per tile]
  DevBuf<uint32_t> d_moves;     // compact traceback: 2-bit move strings, [word][slot]
  DevBuf<uint64_t> d_ctotal;    // compact traceback: alignment bytes used by each chunk
  DevBuf<uint32_t> d_order, d_start, d_alnlen, d_D, d_rerun;
  DevBuf<int32_t> d_status;
  DevBuf<int> d_flags;
  DevBuf<double> d_f, d_corner, d_matd, d_Hout, d_hcol_d;
  DevBuf<int> d_mati, d_hcol_i;
  DevBuf<int> d_band;           // band pipeline kernel: reduction cells, progress words, band rows
  DevBuf<uint32_t> d_ckpt;      // checkpointed traceback: state of the walk between band groups
  DevBuf<uint32_t> d_pack;      // packed sequences for the strip kernel
  DevBuf<uint32_t> d_wz;        // warp kernel: [8 counters][fail list][z bits 2*nzw per slot]
  DevBuf<uint32_t> d_q2raw, d_t2raw;  // packed input as it crossed the ABI (pair-major words)
  DevBuf<uint64_t> d_qwoff, d_twoff;  // ragged packed input: word offsets (count + 1)
  DevBuf<uint32_t> d_mvout;     // compact results: dense move words in pair order
  DevBuf<uint64_t> d_mvoff;     // [count] word offset of every pair's move string
  DevBuf<uint64_t> d_mvbase;    // [chunks + 1] move words before each chunk (running total)
  int unpack_chunk(const Chunk &c);      // packed input -> byte codes of one chunk (uniform batches)
  int unpack_all();                      // ... of the whole batch
  int moves_from_strings();              // !moves_direct: scan + strings -> dense move words
  int download_moves(alnr_batch_out *out);
  uint64_t rerun_total = 0;
  bool profile = false;                 // record events around every fill launch
  std::vector<cudaEvent_t> fill_events; // pairs (begin, end), filled when profile is set
  size_t fill_events_used = 0;
  float last_fill_ms = 0.f;
  int fill_mark(bool end);              // record the next profile event on `stream`
  int collect_fill_ms();                // after a sync: sum of the fill intervals

  ~alnr_batch_impl();
  int init(const alnr_params *p, const alnr_batch *b, bool want_mats);
  int run();                     // enqueue all kernels on `stream`
  int run_chunk(size_t ci);      // phase 1 + phase 2 of one chunk with the cross-stream dependencies
  // host buffers in -> host buffers out, chunks pipelined over three streams
  int stream_all(const alnr_batch *b, alnr_batch_out *out);
  bool resident = true;          // inputs were uploaded by init()
  uint32_t max_chunk = 1u << 22;
  bool ramp = false;             // streaming: short first and last chunks
  int download(alnr_batch_out *out);
  int download_scores(double *f, double *corner, uint32_t *end_yx, int32_t *status);
  int classify();
  DLayout layout_for(const Chunk &c) const;
  DLayout layout_dims(uint32_t m, uint32_t n, bool wave) const;
  bool chunk_compact(const Chunk &c) const;  // alignments of the chunk are written back to back
  bool use_warp(uint32_t m) const;  // warp-per-pair kernel instead of one pair per thread
  int plan_chunks(const uint64_t *q_off, const uint64_t *t_off);
  int plan_and_alloc(const alnr_batch *b);  // chunks + every per-chunk scratch buffer
  double budget_scale = 1.0;                // halved after an out-of-memory, then re-planned
};

}  // namespace alnr
