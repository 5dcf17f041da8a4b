/*
 * aligner_b200.h — C ABI of the B200-native replacement for aligner-core's pairwise
 * alignment hot path.  Plain pointers and sizes only; no C++ or torch types.
 *
 * The reference has no FFI: its boundary is the Rust trait
 *   AlignerTrait::perform_alignment(&mut self, del: f64, ext: f64, matrix: &Array2<f64>,
 *                                   heuristics: Option<Heuristics>)
 *       -> Result<AlignmentResult<T, A>>                 aligner-core/src/lib.rs:27-40
 * implemented by SimpleGlobalAligner / SimpleLocalAligner (aligner-core/src/simple/mod.rs:42,
 * :168) and HeuristicAligner (aligner-core/src/heuristic/mod.rs:36).  Each entry point below
 * names the reference item it replaces; INTEGRATION.md shows the Rust `-sys` binding.
 *
 * Conventions
 *  - Symbols are the reference's numeric codes (aligner-core/src/enums.rs:55-84, :138-147):
 *    Protein A R N D C Q E G H I L K M F P S T W Y V B J Z X = 0..23, DNA A T C G = 0..3,
 *    Blank = 98 in returned alignments.  query has n symbols, target has m symbols.
 *  - matrix is row-major V x V doubles, indexed matrix[target_symbol * V + query_symbol]
 *    (simple/mod.rs:85, :198).
 *  - Directions: 0 Top, 1 Left, 2 Diagonal, 3 Beginning (enums.rs:9-15).
 *  - coords[4] = {q_from, q_to, t_from, t_to} = Alignment.coords ((a,b),(c,d))
 *    (aligner-core/src/alignment.rs:8).
 *  - Every function returns an alnr_status; nothing throws or aborts across the boundary.
 *    Inputs on which the reference panics return a status instead.
 *  - All compute runs on the GPU.  There is no CPU fallback: without a usable CUDA device
 *    every compute entry point returns ALNR_CUDA_ERROR.
 *  - Thread-safe: calls may be issued concurrently from many host threads (the reference's
 *    callers do so, aligner-core/src/statistics/mod.rs:255-286); each call/handle uses its
 *    own CUDA stream.
 */
#ifndef ALIGNER_B200_H
#define ALIGNER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 0 OK, then aligner_core::Error in declaration order (lib.rs:47-57), then
 * aligner_helpers::Error (aligner-helpers/src/lib.rs:12-14), then panics-turned-statuses. */
typedef enum {
  ALNR_OK = 0,
  ALNR_PROTEIN_NOT_FOUND = 1,
  ALNR_CHAR_IS_NOT_MATCHABLE = 2,
  ALNR_UNNECESSARY_ARGUMENT = 3,
  ALNR_MISSING_ARGUMENT = 4,
  ALNR_RESULT_IS_EMPTY = 5,
  ALNR_CALCULATION_ERROR = 6,
  ALNR_VALIDATION_ERROR = 7,
  ALNR_MATRIX_SHAPE_ERROR = 8,
  ALNR_WRONG_MATRIX = 9,         /* transform_matrix: Roots::No (matrices/mod.rs:45) */
  ALNR_EMPTY_INPUT = 10,         /* simple/mod.rs:103-104 unwrap() on an empty sequence */
  ALNR_NO_POSITIVE_CELL = 11,    /* simple/mod.rs:214 underflow: arg-max is (0,0) */
  ALNR_SYMBOL_OUT_OF_RANGE = 12, /* matrix[[t,q]] out of bounds (Blank/Pos/Any in input) */
  ALNR_CUDA_ERROR = 13,
  ALNR_INVALID_ARGUMENT = 14,
  ALNR_OUT_OF_MEMORY = 15
} alnr_status;

enum { ALNR_GLOBAL = 0, ALNR_LOCAL = 1 };
enum { ALNR_DIR_TOP = 0, ALNR_DIR_LEFT = 1, ALNR_DIR_DIAGONAL = 2, ALNR_DIR_BEGINNING = 3 };
#define ALNR_BLANK 98
#define ALNR_POS 99

/* Which arithmetic a call ran in (reported, never chosen by the caller unless forced). */
enum {
  ALNR_NUM_AUTO = 0, /* integer kernels when matrix, del, ext are integral and bounded */
  ALNR_NUM_I16X2 = 1,
  ALNR_NUM_I32 = 2,
  ALNR_NUM_F64 = 3
};

typedef struct {
  int mode;            /* ALNR_GLOBAL: simple/mod.rs:42-144; ALNR_LOCAL: simple/mod.rs:168-264 */
  int volume;          /* V = BioData::volume(): 24 Protein, 4 DNA (enums.rs:398-400, :562-564) */
  double del;          /* perform_alignment(del, ..) */
  double ext;          /* perform_alignment(.., ext, ..) — the reference's stateful rule, not Gotoh */
  const double *matrix; /* V*V, [t][q] */
  int device;          /* CUDA ordinal; -1 = device 0 */
  int numeric;         /* ALNR_NUM_AUTO or a forced domain (tests) */
} alnr_params;

/* ---- library / device ---------------------------------------------------------------- */
const char *alnr_strerror(int status);
const char *alnr_last_error_detail(void); /* thread-local text of the last CUDA failure */
int alnr_device_count(void);              /* 0 when no usable CUDA device */
const char *alnr_build_info(void);        /* "sm_100a ..." */
/* Number of kernel launches issued by this library since load (all threads). */
uint64_t alnr_kernel_launch_count(void);

/* ---- one pair: AlignerTrait::perform_alignment for Simple{Global,Local}Aligner -------- */
typedef struct {
  double f;            /* Alignment.f: 0.0 in global mode (simple/mod.rs:139), max(H) in local */
  double corner_score; /* alignment_matrix[[m, n]] — where the global score lives */
  uint64_t coords[4];
  uint8_t *q_aln;      /* Alignment.query, aln_len codes, library-allocated */
  uint8_t *t_aln;      /* Alignment.target */
  uint64_t aln_len;
  double *H;           /* AlignmentResult.alignment_matrix, (m+1)*(n+1) row-major, if requested */
  uint8_t *D;          /* AlignmentResult.direction_matrix, same shape, 1 byte per cell */
  int numeric;         /* domain actually used */
} alnr_pair_result;

/* want_matrices != 0 also returns H and D (alignment_result.rs:8-9). */
int alnr_align_pair(const alnr_params *p, const uint8_t *q, uint64_t n, const uint8_t *t,
                    uint64_t m, int want_matrices, alnr_pair_result *out);
void alnr_free_pair_result(alnr_pair_result *r);

/* ---- batches of independent pairs (replaces the callers' thread pools:
 *      statistics/mod.rs:255-286, latent-repeat-search/engine/calc.rs:47-69) -------------- */
typedef struct {
  const uint8_t *q;      /* concatenated query codes */
  const uint64_t *q_off; /* count+1 offsets */
  const uint8_t *t;      /* concatenated target codes */
  const uint64_t *t_off; /* count+1 offsets */
  uint64_t count;
  /* Opt-in compact input (alnr_align_batch / alnr_batch_create* / alnr_score_batch only; V == 4):
   * 2-bit DNA codes, 16 per 32-bit word, code k of a sequence at bits [2 * (k % 16), +2) of its word
   * k / 16; every sequence starts on a word boundary, sequences densely one after the other (sequence
   * i starts at word sum_{j<i} ceil(len_j / 16)); unused high bits of a last word are ignored.  Set
   * both or neither; when set, q / t are ignored and may be NULL.  q_off / t_off keep their meaning
   * (symbol offsets, i.e. the lengths).  A binding that walks Vec<DNA> (aligner-core/src/enums.rs:
   * 138-147) to produce byte codes can pack 2 bits instead: 4x fewer bytes cross PCIe. */
  const uint32_t *q2;
  const uint32_t *t2;
} alnr_batch;

/* Caller-provided output arrays (host memory; pinned memory makes copies faster).
 * Any pointer may be NULL to skip that output.  Alignment i occupies
 * [aln_off[i], aln_off[i] + aln_len[i]) of q_aln / t_aln; aln_off is authoritative.  q_aln/t_aln
 * need q_off[count] + t_off[count] + count bytes (the reference's worst case, n_i + m_i + 1 per
 * pair); bytes outside the alignments are unspecified.  Batches of equal-length short pairs are
 * written back to back per chunk (only the bytes that exist cross PCIe); other batches put
 * alignment i right-aligned into its own slot starting at q_off[i] + t_off[i] + i. */
typedef struct {
  double *f;           /* [count] */
  double *corner;      /* [count] */
  uint64_t *coords;    /* [count*4] */
  uint64_t *aln_off;   /* [count] */
  uint32_t *aln_len;   /* [count] */
  uint8_t *q_aln;
  uint8_t *t_aln;
  int32_t *status;     /* [count] per-pair status (EMPTY_INPUT, NO_POSITIVE_CELL, ...) */
  /* Opt-in compact result format: set `moves` (then q_aln / t_aln / aln_off must be NULL).  The
   * alignment of pair i is returned as the walk the reference's traceback takes
   * (simple/mod.rs:107-127, :220-242): start cell end_yx[2i], end_yx[2i+1] = 1-based (target,
   * query) index of the LAST aligned column (the arg-max in local mode, (m, n) in global mode) and
   * aln_len[i] - 1 moves, move k = Direction of the k-th step away from that cell (0 Top, 1 Left,
   * 2 Diagonal), 2 bits per move, 16 per word, move k at bits [2 * (k % 16), +2) of word
   * moves_off[i] + k / 16.  Move strings are dense: moves_off[i] = sum_{j<i} ceil((aln_len[j] - 1)
   * / 16) (pairs with a status own no words).  `moves` needs alnr_moves_capacity() words.
   * Alignment.query / .target are private to aligner-core (alignment.rs:6-7): callers see coords,
   * f, get_frequency_matrix() and get_alignment() — all of which alnr_expand_alignments() (or a
   * lazily expanding binding) derives from this record; ~70 B instead of ~390 B per 150 bp pair. */
  uint32_t *end_yx;    /* [count*2]; also available in the string format */
  uint64_t *moves_off; /* [count], optional */
  uint32_t *moves;
} alnr_batch_out;

/* Words `moves` must hold for this batch (worst case ceil((n_i + m_i) / 16) per pair). */
uint64_t alnr_moves_capacity(const alnr_batch *b);

/* Host-side expansion of the compact format (threads <= 0: all host cores).  For the pairs
 * [first, first + count) of batch b (byte or 2-bit inputs) it writes, indexed by i - first: coords,
 * aln_len, aln_off (alignments back to back from 0, needs sum of aln_len bytes; when the strings
 * are NULL only coords / aln_off are produced), q_aln, t_aln exactly as the string format would
 * (the seeded last column included, simple/mod.rs:102-105), and optionally the V x V frequency
 * matrices of Alignment::get_frequency_matrix (alignment.rs:13-23), freq[(i - first) * V * V ...].
 * compact->moves_off may be NULL only when first == 0 (offsets are then re-derived). */
int alnr_expand_alignments(int mode, int volume, const alnr_batch *b, const alnr_batch_out *compact,
                           uint64_t first, uint64_t count, alnr_batch_out *expanded, double *freq,
                           int threads);

/* Host buffers in, host buffers out (H2D + kernels + D2H, chunked and overlapped). */
int alnr_align_batch(const alnr_params *p, const alnr_batch *b, alnr_batch_out *out);

/* The same call over the GPUs of one box (BASELINE north_star (3); replaces the Kafka fan-out of
 * aligner-web/src/bin/dispatcher/handlers.rs:193-251 + node/main.rs:61-74 for independent pairs):
 * the batch is cut into n_gpus contiguous shards of pairs balanced by the sum of n * m, every shard
 * runs on its own device (shard k on device k % alnr_device_count(), p->device is ignored) from its
 * own host thread, and writes straight into the caller's arrays at its offset.  No collective.
 * n_gpus <= 0: one shard per device.  Results are those of alnr_align_batch for every n_gpus, except
 * that in the compact format the move strings are dense per SHARD (moves_off is mandatory and
 * authoritative; shard k's words start at the worst-case capacity of the shards before it). */
int alnr_align_batch_multi(const alnr_params *p, const alnr_batch *b, alnr_batch_out *out, int n_gpus);

/* Device-resident form of the same call: upload once, run many times, download. */
typedef struct alnr_batch_handle alnr_batch_handle;
int alnr_batch_create(const alnr_params *p, const alnr_batch *b, alnr_batch_handle **h);
/* Same, but the traceback stops at the 2-bit move strings (compact result format, see
 * alnr_batch_out.moves); read with alnr_batch_download and `moves` set. */
int alnr_batch_create_compact(const alnr_params *p, const alnr_batch *b, alnr_batch_handle **h);
int alnr_batch_run(alnr_batch_handle *h);               /* kernels only, asynchronous */
int alnr_batch_sync(alnr_batch_handle *h);
int alnr_batch_download(alnr_batch_handle *h, alnr_batch_out *out);
void *alnr_batch_stream(alnr_batch_handle *h);          /* cudaStream_t the kernels run on */
/* Runs the batch `steps` times back to back between two CUDA events recorded on its own
 * stream and waits; *ms_total = device time of all steps.  If fill_ms_total != NULL every
 * fill-kernel launch is additionally bracketed by events and their sum is returned. */
int alnr_batch_run_timed(alnr_batch_handle *h, int steps, float *ms_total, float *fill_ms_total);
int alnr_batch_numeric(alnr_batch_handle *h);           /* domain selected for this batch */
uint64_t alnr_batch_rerun_count(alnr_batch_handle *h);  /* pairs whose speculation was redone */
void alnr_batch_destroy(alnr_batch_handle *h);

/* ---- score-only batches ------------------------------------------------------------------------
 * The p-value estimate (aligner-core/src/statistics/mod.rs:240-307) aligns one query against
 * 4999 shuffled copies of the target and keeps ONLY `alignment.f` of each call (:268-276).  These
 * entry points run the same fill kernels without direction scratch or traceback:
 *   f[i]       Alignment.f (local: max H; global: 0.0, simple/mod.rs:139)
 *   corner[i]  alignment_matrix[[m, n]] (the global score)
 *   end_yx[2i], end_yx[2i+1]  1-based (target, query) cell the traceback would start from: the
 *              first maximum in row-major order (simple/mod.rs:212), or (m, n) in global mode
 *   status[i]  as in alnr_batch_out
 * Any output pointer may be NULL. */
int alnr_score_batch(const alnr_params *p, const alnr_batch *b, double *f, double *corner,
                     uint32_t *end_yx, int32_t *status);
/* Device-resident form: a handle for alnr_batch_run / _run_timed / _sync; read the results with
 * alnr_batch_download_scores (alnr_batch_download accepts f / corner / status only). */
int alnr_batch_create_scores(const alnr_params *p, const alnr_batch *b, alnr_batch_handle **h);
int alnr_batch_download_scores(alnr_batch_handle *h, double *f, double *corner, uint32_t *end_yx,
                               int32_t *status);

/* ---- score-only batch that allocates nothing ----------------------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same results as alnr_score_batch (f, corner, end_yx, status — see above) from ONE kernel
 * launch on `stream` (a cudaStream_t, NULL = the default stream): no cudaMalloc / cudaFree, no
 * copies, no pool.  The call returns after enqueueing; synchronise the stream before reading.
 *  - d_q, d_q_off, d_t, d_t_off, d_f, d_corner, d_end_yx, d_status are DEVICE pointers on device
 *    p->device (-1 = device 0); output pointers may be NULL.  p->matrix (host, V*V) goes by value.
 *  - max_query_len / max_target_len: the caller's promise about the longest sequences.  A longer
 *    pair gets status ALNR_INVALID_ARGUMENT (it is not computed).
 *  - The scratch (one score column per pair) is a static device array of
 *    ALNR_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Limit:
 *      ceil(count / 32) * 32 * (max_target_len + 1) * (4 integer problems | 8 FP64) <= that size
 *    (e.g. 8192 pairs x 500 rows integer, 4096 pairs x 500 rows FP64); beyond it the call returns
 *    ALNR_INVALID_ARGUMENT with the arithmetic in alnr_last_error_detail().  Split the batch.
 *  - Every call on a device shares the array and it keeps the last call's contents (nothing in it
 *    needs clearing: each pair writes its column 0 first).  Calls that overlap on the same GPU —
 *    different streams or host threads — RACE on it: serialise them. */
#define ALNR_NOALLOC_SCRATCH_BYTES (16u << 20)
int alnr_score_batch_noalloc(const alnr_params *p, const uint8_t *d_q, const uint64_t *d_q_off,
                             const uint8_t *d_t, const uint64_t *d_t_off, uint64_t count,
                             uint32_t max_query_len, uint32_t max_target_len, double *d_f,
                             double *d_corner, uint32_t *d_end_yx, int32_t *d_status, void *stream);
uint64_t alnr_noalloc_scratch_bytes(void);

/* ---- PWM aligner --------------------------------------------------------
 * PWMAligner::perform_alignment (aligner-core/src/pwm/mod.rs:29-126): Smith-Waterman-like local
 * alignment of a DNA window (the DP rows) against the L columns of a 4 x L position-weight matrix,
 * score = matrix[[symbol, x-1]], same tie / Beginning / stateful-penalty rules as the local
 * aligner.  Differences the reference has and this path keeps: no seed pair before the traceback
 * (:81-104) and no error when the arg-max is (0,0) — the alignment is empty, f = 0, coords
 * ((1,1),(1,1)).  The reference's batch workload aligns sliding windows of one chromosome to one
 * PWM on OS threads (bin/latent-repeat-search/engine/calc.rs:47-69, :107-137): here the sequence
 * is uploaded once (alnr_pwm_seq) and every call aligns any set of its windows to one PWM. */
typedef struct {
  double del, ext;
  const double *pwm;  /* rows x length, row-major: matrix[[symbol, position]] */
  uint32_t rows;      /* must be 4, else ALNR_MATRIX_SHAPE_ERROR (pwm/mod.rs:40-42) */
  uint32_t length;    /* L = matrix.shape()[1] */
  int device;         /* must be the device of the sequence; -1 = that device */
  int numeric;        /* ALNR_NUM_AUTO, or ALNR_NUM_I32 / ALNR_NUM_F64 forced (tests) */
} alnr_pwm_params;

typedef struct alnr_pwm_seq alnr_pwm_seq; /* DNA codes 0..3 resident in HBM */
int alnr_pwm_seq_create(const uint8_t *seq, uint64_t len, int device, alnr_pwm_seq **h);
uint64_t alnr_pwm_seq_len(const alnr_pwm_seq *h);
void alnr_pwm_seq_destroy(alnr_pwm_seq *h);

/* Window i = seq[start_i, start_i + len_i).  Explicit: starts / lens (both non-NULL).  Strided
 * (starts == NULL): start_i = first + i * step, len_i = min(window, seq_len - start_i) — the
 * j..border windows of calc.rs:50-56 with step = query_offset, window = repeat_length + query_offset. */
typedef struct {
  uint64_t count;
  const uint64_t *starts;
  const uint32_t *lens;
  uint64_t first, step;
  uint32_t window;
} alnr_windows;

/* Caller-provided host arrays; any pointer may be NULL.  With numbered == query == freq == NULL the
 * call is score-only (no direction scratch, no traceback).  Window i owns entries
 * [i * aln_slot, (i + 1) * aln_slot) of numbered / query; its alignment is the aln_len[i] entries
 * from aln_off[i] on (right-aligned in the slot).  aln_slot >= L + longest window. */
typedef struct {
  double *f;          /* [count] PWMAlignment.f = max of the matrix */
  uint64_t *coords;   /* [count*4] ((x_from, x_to), (y_from, y_to)), pwm/mod.rs:118-121 */
  int32_t *status;    /* [count] 0 or ALNR_SYMBOL_OUT_OF_RANGE */
  uint64_t *aln_off;  /* [count] */
  uint32_t *aln_len;  /* [count] */
  uint32_t *numbered; /* PWMAlignment.numbered: profile position 1..L, 0 = gap (Top move) */
  uint8_t *query;     /* PWMAlignment.query: window symbol, ALNR_BLANK = gap (Left move) */
  uint64_t aln_slot;
  double *freq;       /* [count*4*L] PWMAlignment::get_frequency_matrix (alignment.rs:55-65) */
  double *H;          /* count == 1 only: alignment_matrix, (m+1) x (L+1) row-major */
  uint8_t *D;         /* count == 1 only: direction_matrix, 1 byte per cell */
} alnr_pwm_out;

int alnr_pwm_align_windows(const alnr_pwm_params *p, alnr_pwm_seq *seq, const alnr_windows *w,
                           alnr_pwm_out *out);
int alnr_pwm_last_numeric(void); /* numeric domain of this thread's last alnr_pwm_align_windows */

/* ---- heuristic matrix search ---------------------------------------------------------------
 * transform_matrix (aligner-helpers/src/matrices/mod.rs:19-68): rows x cols matrix, frequencies
 * [rows]; out may alias nothing.  Returns ALNR_WRONG_MATRIX on Roots::No. */
int alnr_transform_matrix(const double *matrix, int rows, int cols, double kd, double r_squared,
                          const double *frequencies, double *out);

/* Items of a heuristic batch: item i aligns pair pair_idx[i] starting from candidate matrix
 * mat_idx[i].  pair_idx == NULL means the full grid, item = pair * n_matrices + k. */
typedef struct {
  const uint32_t *pair_idx;
  const uint32_t *mat_idx;
  uint64_t count;            /* items (ignored for the full grid) */
  const double *matrices;    /* n_matrices * V * V, each [t][q] */
  uint32_t n_matrices;
  double kd;                 /* Heuristics.kd          (lib.rs:21-25) */
  double r_squared;          /* Heuristics.r_squared; 0 => V*V (heuristic/mod.rs:48-50) */
  const double *frequencies; /* Heuristics.frequencies per PAIR: count_pairs * V */
} alnr_heuristic_items;

typedef struct {          /* caller-provided, any pointer may be NULL */
  double *f;              /* [items] Alignment.f of the LAST pass (heuristic/mod.rs:73-76) */
  int32_t *passes;        /* [items] local alignments run (>= 2 when status == 0) */
  int32_t *status;        /* [items] */
  uint64_t *coords;       /* [items*4] of the last pass */
  double *matrices;       /* [items*V*V] AlignmentResult.matrix = Some(T): the matrix of the last pass */
} alnr_heuristic_out;

/* HeuristicAligner::perform_alignment (aligner-core/src/heuristic/mod.rs:36-78) over many
 * (pair, matrix) items; p->matrix is ignored, p->mode must be ALNR_LOCAL. */
int alnr_heuristic_batch(const alnr_params *p, const alnr_batch *b, const alnr_heuristic_items *items,
                         alnr_heuristic_out *out);

/* The outer search the web layer runs through Kafka + MariaDB (dispatcher/handlers.rs:193-251,
 * node/main.rs:61-74): all pairs x all candidate matrices.  fitness[k] = sum_pairs f(pair, k)
 * (fixed blocked order, see alnr_fitness_block_sums), best_k/best_f per pair (first maximum); f_grid/status_grid optional
 * [pairs * n_matrices]. */
int alnr_matrix_search(const alnr_params *p, const alnr_batch *b, const double *matrices,
                       uint32_t n_matrices, double kd, double r_squared, const double *frequencies,
                       double *fitness, uint32_t *best_k, double *best_f, double *f_grid,
                       int32_t *status_grid);

/* alnr_matrix_search over n_gpus devices of one box: pairs sharded on ALNR_FITNESS_BLOCK boundaries
 * (balanced by the sum of n * m), one host thread and stream set per device, the f grid written in
 * place, fitness reduced in the fixed blocked order below — bit-identical for every n_gpus.  (One
 * process per GPU exchanges the block sums with one NCCL all-reduce instead: aligner_b200/sharding.py.) */
int alnr_matrix_search_multi(const alnr_params *p, const alnr_batch *b, const double *matrices,
                             uint32_t n_matrices, double kd, double r_squared, const double *frequencies,
                             double *fitness, uint32_t *best_k, double *best_f, double *f_grid,
                             int32_t *status_grid, int n_gpus);

/* Fixed-order fitness reduction, the one step of the search that crosses GPUs.
 * block_sums[b * n_matrices + k] = sum of f_grid[i * n_matrices + k] over the pairs i of block b
 * (ALNR_FITNESS_BLOCK consecutive pairs, added in pair order; status != 0 items skipped).
 * fitness[k] = sum over blocks in block order.  alnr_matrix_search uses exactly this order, so a
 * search sharded over N GPUs on block boundaries (blocks exchanged with one NCCL all-reduce,
 * every block owned by one rank) gives bit-identical fitness for every N. */
#define ALNR_FITNESS_BLOCK 256
int alnr_fitness_block_sums(const double *f_grid, const int32_t *status_grid, uint64_t n_pairs,
                            uint32_t n_matrices, double *block_sums);
int alnr_fitness_from_blocks(const double *block_sums, uint64_t n_blocks, uint32_t n_matrices,
                             double *fitness);

/* Roofline denominators measured on the device: arithmetic INT32 ops/s (an op = one
 * 32-bit add / max / logic op / multiply-add lane) of alu-pipe-only chains and of a 1:1
 * alu-pipe + fma-pipe mix over all SMs. */
int alnr_measure_int_peak(int device, double *alu_ops_per_s, double *mixed_ops_per_s);
/* FP64 denominators for the heuristic matrix search (non-integral matrices force doubles):
 * ops/s of independent DADD chains, and of DADD + fmax chains (an op = one f64 add or max). */
int alnr_measure_fp64_peak(int device, double *add_ops_per_s, double *addmax_ops_per_s);

/* pinned host memory helpers for callers without a CUDA binding */
void *alnr_host_alloc(uint64_t bytes);
void alnr_host_free(void *p);
/* Device buffers are cached between calls (cudaMalloc/cudaFree are slow and synchronising);
 * this returns every cached buffer to the driver. */
void alnr_release_cached_memory(void);

#ifdef __cplusplus
}
#endif
#endif
