// multi.cu — the batch partitioning across the GPUs of one box, behind the C ABI.
//
// The reference fans a search out as one Kafka message per (pair, candidate matrix) to stateless
// workers and collects the results in MariaDB (aligner-web/src/bin/dispatcher/handlers.rs:193-251,
// aligner-web/src/bin/node/main.rs:61-74).  Here one call takes the whole batch, cuts it into one
// contiguous shard of pairs per GPU (balanced by the sum of n * m) and drives every device from its
// own host thread with its own streams; results land directly in the caller's arrays at the shard's
// offset.  There is no data-path collective: pairs are independent.  The matrix search needs one
// reduction — fitness[k] = sum over pairs of f(pair, k) — and keeps the FIXED blocked order of
// alnr_fitness_block_sums: shards start on block boundaries and the block sums are added in block
// order on the host, so the fitness bits do not depend on the number of GPUs.
// (One process per GPU with an NCCL all-reduce of the block sums is the alternative shape:
// aligner_b200/sharding.py.)
#include <algorithm>
#include <cmath>
#include <thread>
#include <vector>

#include "host_util.h"

namespace {

// Shards [lo_k, lo_{k+1}) of `count` items with weights w(i), boundaries rounded to `align`.
template <typename W>
std::vector<uint64_t> shard_bounds(uint64_t count, int shards, uint64_t align, W weight) {
  std::vector<uint64_t> b((size_t)shards + 1, count);
  b[0] = 0;
  if (shards <= 1 || count == 0) return b;
  double total = 0;
  for (uint64_t i = 0; i < count; ++i) total += weight(i);
  double acc = 0;
  int k = 1;
  for (uint64_t i = 0; i < count && k < shards; ++i) {
    acc += weight(i);
    while (k < shards && acc >= total * k / shards) {
      uint64_t cut = (i + 1 + align / 2) / align * align;
      cut = std::min<uint64_t>(std::max<uint64_t>(cut, b[k - 1]), count);
      b[k++] = cut;
    }
  }
  return b;
}

// A shard of a batch as a batch of its own: offsets rebased to 0, symbol pointers moved.
struct SubBatch {
  std::vector<uint64_t> qo, to;
  alnr_batch b{};
  uint64_t aln_base = 0;  // first byte of the shard's region in q_aln / t_aln (string format)
  uint64_t mv_base = 0;   // first word of the shard's region in `moves`
};

void make_sub(const alnr_batch *src, uint64_t lo, uint64_t hi, uint64_t q2w, uint64_t t2w, SubBatch &s) {
  const uint64_t c = hi - lo;
  s.qo.resize(c + 1);
  s.to.resize(c + 1);
  const uint64_t q0 = src->q_off[lo], t0 = src->t_off[lo];
  for (uint64_t i = 0; i <= c; ++i) { s.qo[i] = src->q_off[lo + i] - q0; s.to[i] = src->t_off[lo + i] - t0; }
  s.b.q = src->q ? src->q + q0 : nullptr;
  s.b.t = src->t ? src->t + t0 : nullptr;
  s.b.q_off = s.qo.data();
  s.b.t_off = s.to.data();
  s.b.count = c;
  s.b.q2 = src->q2 ? src->q2 + q2w : nullptr;
  s.b.t2 = src->t2 ? src->t2 + t2w : nullptr;
  s.aln_base = q0 + t0 + lo;
}

int resolve_devices(int n_gpus, int &shards, int &ndev) {
  ndev = alnr_device_count();
  if (ndev < 1) {
    alnr::set_error_detail("no usable CUDA device (this library has no CPU fallback)");
    return ALNR_CUDA_ERROR;
  }
  shards = n_gpus <= 0 ? ndev : n_gpus;  // more shards than devices: shard k runs on device k % ndev
  if (shards > 64) return ALNR_INVALID_ARGUMENT;
  return ALNR_OK;
}

}  // namespace

extern "C" int alnr_align_batch_multi(const alnr_params *p, const alnr_batch *b, alnr_batch_out *out, int n_gpus) {
  if (!p || !b || !out) return ALNR_INVALID_ARGUMENT;
  if (b->count && (!b->q_off || !b->t_off)) return ALNR_INVALID_ARGUMENT;
  if (out->moves && !out->moves_off) return ALNR_INVALID_ARGUMENT;  // shards are dense one by one, not globally
  int shards = 1, ndev = 0;
  int st = resolve_devices(n_gpus, shards, ndev);
  if (st) return st;
  const uint64_t P = b->count;
  shards = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)shards, (P + 63) / 64));
  if (shards == 1) {
    alnr_params p1 = *p;
    if (p1.device < 0) p1.device = 0;
    return alnr_align_batch(&p1, b, out);
  }
  try {
    const std::vector<uint64_t> cut = shard_bounds(P, shards, 64, [&](uint64_t i) {
      return (double)(b->q_off[i + 1] - b->q_off[i]) * (double)(b->t_off[i + 1] - b->t_off[i]);
    });
    // word offsets of every shard's first sequence (2-bit inputs) and of its move-string region
    std::vector<uint64_t> q2w((size_t)shards + 1, 0), t2w((size_t)shards + 1, 0), mvw((size_t)shards + 1, 0);
    for (int k = 0; k < shards; ++k) {
      uint64_t a = 0, c = 0, mv = 0;
      for (uint64_t i = cut[k]; i < cut[k + 1]; ++i) {
        const uint64_t n = b->q_off[i + 1] - b->q_off[i], m = b->t_off[i + 1] - b->t_off[i];
        a += (n + 15) >> 4; c += (m + 15) >> 4; mv += (n + m + 15) >> 4;
      }
      q2w[k + 1] = q2w[k] + a; t2w[k + 1] = t2w[k] + c; mvw[k + 1] = mvw[k] + mv;
    }
    std::vector<int> status((size_t)shards, ALNR_OK);
    std::vector<std::thread> th;
    for (int k = 0; k < shards; ++k) {
      th.emplace_back([&, k] {
        const uint64_t lo = cut[k], hi = cut[k + 1];
        if (hi == lo) return;
        SubBatch s;
        make_sub(b, lo, hi, q2w[k], t2w[k], s);
        alnr_params pk = *p;
        pk.device = k % ndev;
        alnr_batch_out o{};
        o.f = out->f ? out->f + lo : nullptr;
        o.corner = out->corner ? out->corner + lo : nullptr;
        o.coords = out->coords ? out->coords + 4 * lo : nullptr;
        o.aln_off = out->aln_off ? out->aln_off + lo : nullptr;
        o.aln_len = out->aln_len ? out->aln_len + lo : nullptr;
        o.q_aln = out->q_aln ? out->q_aln + s.aln_base : nullptr;
        o.t_aln = out->t_aln ? out->t_aln + s.aln_base : nullptr;
        o.status = out->status ? out->status + lo : nullptr;
        o.end_yx = out->end_yx ? out->end_yx + 2 * lo : nullptr;
        o.moves_off = out->moves_off ? out->moves_off + lo : nullptr;
        o.moves = out->moves ? out->moves + mvw[k] : nullptr;
        int r = alnr_align_batch(&pk, &s.b, &o);
        if (r == ALNR_OK) {  // offsets are relative to the shard's regions: make them the caller's
          if (o.aln_off) for (uint64_t i = 0; i < hi - lo; ++i) o.aln_off[i] += s.aln_base;
          if (o.moves_off) for (uint64_t i = 0; i < hi - lo; ++i) o.moves_off[i] += mvw[k];
        }
        status[k] = r;
      });
    }
    for (auto &t : th) t.join();
    for (int r : status) if (r) return r;
  } catch (...) {
    return ALNR_OUT_OF_MEMORY;
  }
  return ALNR_OK;
}

extern "C" int alnr_matrix_search_multi(const alnr_params *p, const alnr_batch *b, const double *matrices,
                                        uint32_t n_matrices, double kd, double r_squared,
                                        const double *frequencies, double *fitness, uint32_t *best_k,
                                        double *best_f, double *f_grid, int32_t *status_grid, int n_gpus) {
  if (!p || !b || !fitness || !matrices || !frequencies || n_matrices == 0) return ALNR_INVALID_ARGUMENT;
  if (b->count && (!b->q || !b->t || !b->q_off || !b->t_off)) return ALNR_INVALID_ARGUMENT;
  int shards = 1, ndev = 0;
  int st = resolve_devices(n_gpus, shards, ndev);
  if (st) return st;
  const uint64_t P = b->count, K = n_matrices;
  const int V = p->volume;
  try {
    std::vector<double> fbuf;
    std::vector<int32_t> sbuf;
    double *fg = f_grid;
    int32_t *sg = status_grid;
    if (!fg) { fbuf.resize(P * K); fg = fbuf.data(); }
    if (!sg) { sbuf.resize(P * K); sg = sbuf.data(); }
    // shards on fitness-block boundaries: every block of ALNR_FITNESS_BLOCK pairs has one owner
    const std::vector<uint64_t> cut = shard_bounds(P, shards, ALNR_FITNESS_BLOCK, [&](uint64_t i) {
      return (double)(b->q_off[i + 1] - b->q_off[i]) * (double)(b->t_off[i + 1] - b->t_off[i]);
    });
    std::vector<int> status((size_t)shards, ALNR_OK);
    std::vector<std::thread> th;
    for (int k = 0; k < shards; ++k) {
      th.emplace_back([&, k] {
        const uint64_t lo = cut[k], hi = cut[k + 1];
        if (hi == lo) return;
        SubBatch s;
        make_sub(b, lo, hi, 0, 0, s);
        alnr_params pk = *p;
        pk.device = k % ndev;
        alnr_heuristic_items it{nullptr, nullptr, 0, matrices, n_matrices, kd, r_squared, frequencies + lo * (uint64_t)V};
        alnr_heuristic_out o{fg + lo * K, nullptr, sg + lo * K, nullptr, nullptr};
        status[k] = alnr_heuristic_batch(&pk, &s.b, &it, &o);
      });
    }
    for (auto &t : th) t.join();
    for (int r : status) if (r) return r;
    // the reduction: block sums in pair order, blocks in block order (identical for every n_gpus)
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
  } catch (...) {
    return ALNR_OUT_OF_MEMORY;
  }
  return ALNR_OK;
}
