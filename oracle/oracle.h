/*
 * oracle.h — CPU restatement of aligner-core's pairwise alignment hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load liboracle.so.  The product (aligner_b200/) never links or calls it.
 *
 * The reference is pure Rust and no Rust toolchain exists in this image, so the real
 * crate cannot be compiled (oracle/_ref is therefore absent).  This is a literal C++
 * restatement of the cited Rust lines, in the reference's own memory layout
 * (full (m+1)x(n+1) f64 score matrix + 1-byte direction matrix, x-outer/y-inner sweep).
 *
 * Pinning: the GLOBAL fill (scores + directions) is pinned by the reference's only
 * golden vector, src/tests/test_alignment.rs:14-39 (tests/golden/legacy_global.json).
 * Local mode, the stateful del/ext penalty, arg-max tie rule, transform_matrix bit
 * patterns and the heuristic loop are pinned by NO reference test: "parity unpinned"
 * for those; they are cross-checked against an independent Python restatement
 * (oracle/py_restatement.py) only.
 *
 * All citations are relative to /root/reference/.
 */
#ifndef ALIGNER_ORACLE_H
#define ALIGNER_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Direction codes = declaration order of aligner-core/src/enums.rs:9-15 */
enum { ORC_TOP = 0, ORC_LEFT = 1, ORC_DIAGONAL = 2, ORC_BEGINNING = 3 };

/* Blank symbol code, enums.rs:81 (Protein::Blank = 98) and :144 (DNA::Blank = 98) */
#define ORC_BLANK 98
#define ORC_POS 99

/* status codes (panics of the reference become statuses) */
enum {
  ORC_OK = 0,
  ORC_EMPTY_INPUT = 1,       /* simple/mod.rs:103-104 unwrap on empty Vec */
  ORC_NO_POSITIVE_CELL = 2,  /* simple/mod.rs:214 usize underflow when argmax == (0,0) */
  ORC_SYMBOL_OUT_OF_RANGE = 3, /* matrix[[t,q]] out of bounds */
  ORC_WRONG_MATRIX = 4,      /* aligner-helpers/src/lib.rs:12-14 WrongMatrixSpecified */
  ORC_BUFFER_TOO_SMALL = 5,
  ORC_MATRIX_SHAPE_ERROR = 6 /* pwm/mod.rs:40-42 MatrixShapeError */
};

/* enums.rs:18-28 and :30-46.  Returns the max through *v, the direction as return value. */
int orc_get_direction(double top, double left, double diag, double *v);
int orc_get_direction_with_beginning(double top, double left, double diag, double *v);

/*
 * One pairwise alignment in the reference's faithful layout.
 *   mode 0 = SimpleGlobalAligner::perform_alignment  simple/mod.rs:42-144
 *   mode 1 = SimpleLocalAligner::perform_alignment   simple/mod.rs:168-264
 * q[n] = query codes, t[m] = target codes, matrix = row-major V x V, indexed [t][q]
 * (simple/mod.rs:85, :198).
 * Outputs (any pointer may be NULL to skip):
 *   H  (m+1)*(n+1) doubles row-major [y][x];  D same shape, 1 byte per cell
 *   q_aln/t_aln: capacity aln_cap bytes each (m+n+2 always suffices); *aln_len
 *   coords[4] = {q_from, q_to, t_from, t_to}  (alignment.rs:8 ((a,b),(c,d)))
 *   *f = Alignment.f ; *corner = H[m][n]
 */
int orc_align_pair(int mode, const uint8_t *q, size_t n, const uint8_t *t, size_t m,
                   double del, double ext, const double *matrix, int V,
                   double *H, uint8_t *D, uint8_t *q_aln, uint8_t *t_aln, size_t aln_cap,
                   size_t *aln_len, uint64_t *coords, double *f, double *corner);

/*
 * Lean layout of the same recurrence (rolling column + 2-bit directions), for
 * 10^5-scale checks: same results, no H.  d_sha_hex (>= 17 bytes) gets the first 16 hex
 * chars of sha256 over the row-major D bytes (one byte per cell, codes as above) when
 * non-NULL.
 */
int orc_align_pair_lean(int mode, const uint8_t *q, size_t n, const uint8_t *t, size_t m,
                        double del, double ext, const double *matrix, int V,
                        uint8_t *q_aln, uint8_t *t_aln, size_t aln_cap, size_t *aln_len,
                        uint64_t *coords, double *f, double *corner, char *d_sha_hex);

/* alignment.rs:13-23: V x V counts, F[t][q] += 1 over columns without Blank. */
void orc_frequency_matrix(const uint8_t *q_aln, const uint8_t *t_aln, size_t len, int V,
                          double *F);
/* alignment.rs:25-42: consensus line. */
void orc_get_alignment(const uint8_t *q_aln, const uint8_t *t_aln, size_t len,
                       const double *matrix, int V, uint8_t *out);

/* aligner-helpers/src/matrices/mod.rs:19-68 (R x C matrix, freq[R]) */
int orc_transform_matrix(const double *matrix, int R, int C, double k_d, double r_squared,
                         const double *freq, double *out);
/* matrices/mod.rs:8-17 */
double orc_get_threshold(int dim);
/* ndarray 0.15.4 sum() order and ndarray-stats 0.5.0 l2_dist (restated from memory) */
double orc_ndarray_sum(const double *x, size_t len);
double orc_l2_dist(const double *a, const double *b, size_t len);
/* roots 0.0.7 find_roots_quadratic(a2,a1,a0): returns number of roots (0,1,2) */
int orc_find_roots_quadratic(double a2, double a1, double a0, double *roots);

/*
 * heuristic/mod.rs:36-78 HeuristicAligner::perform_alignment.
 * Returns the LAST (non-improving) local alignment and the matrix that produced it
 * (out_matrix, V*V).  *passes = number of SW passes run (>= 2).
 */
int orc_heuristic(const uint8_t *q, size_t n, const uint8_t *t, size_t m, double del,
                  double ext, const double *matrix, int V, double kd, double r_squared,
                  const double *freq, uint8_t *q_aln, uint8_t *t_aln, size_t aln_cap,
                  size_t *aln_len, uint64_t *coords, double *f, double *out_matrix,
                  int *passes);

/*
 * Batch driver used ONLY as the timed CPU baseline and as the parity checker for
 * batched GPU calls: pairs are concatenated codes + uint64 offsets (count+1 entries).
 * Faithful layout per pair (allocations included, as the reference allocates per call),
 * std::thread over pairs, `threads` workers.
 * Outputs per pair: f, corner, coords[4], aln_len; aligned strings into slots of
 * `slot` bytes each (q_aln_all + i*slot), written left-aligned.
 */
int orc_align_batch(int mode, const uint8_t *qs, const uint64_t *qoff, const uint8_t *ts,
                    const uint64_t *toff, size_t count, double del, double ext,
                    const double *matrix, int V, int threads, double *f, double *corner,
                    uint64_t *coords, uint64_t *aln_len, uint8_t *q_aln_all,
                    uint8_t *t_aln_all, size_t slot, int *status);

/* heuristic over items (pair i, matrix k): item = (pair_idx[i], mat_idx[i]) */
int orc_heuristic_batch(const uint8_t *qs, const uint64_t *qoff, const uint8_t *ts,
                        const uint64_t *toff, const uint32_t *pair_idx,
                        const uint32_t *mat_idx, size_t items, double del, double ext,
                        const double *matrices, int V, double kd, double r_squared,
                        const double *freqs /* per pair, V each */, int threads, double *f,
                        int *passes, double *out_matrices /* items*V*V or NULL */,
                        int *status);

/*
 * pwm/mod.rs:29-126 PWMAligner::perform_alignment (oracle_pwm.cpp; parity unpinned, see there).
 * query: m window symbols (rows); pwm: rows x L row-major, rows must be 4.
 * H_out / D_out: optional (m+1) x (L+1).  numbered / q_aln: PWMAlignment.numbered / .query
 * (0 = gap in the profile, ORC_BLANK = gap in the window), *aln_len entries.
 * coords[4] = {x_from, x_to, y_from, y_to} (pwm/mod.rs:118-121).
 */
int orc_pwm_align(const uint8_t *query, size_t m, const double *pwm, size_t rows, size_t L,
                  double del, double ext, double *H_out, uint8_t *D_out, uint32_t *numbered,
                  uint8_t *q_aln, size_t aln_cap, size_t *aln_len, uint64_t *coords, double *f);
/* alignment.rs:55-65 PWMAlignment::get_frequency_matrix -> out[4 * L] */
void orc_pwm_frequency_matrix(const uint32_t *numbered, const uint8_t *q_aln, size_t len, size_t L,
                              double *out);
/* windows seq[starts[i], +lens[i]) against one PWM on `threads` threads
 * (latent-repeat-search/engine/calc.rs:47-69, :107-137); alignment slots of `slot` entries. */
void orc_pwm_windows(const uint8_t *seq, const uint64_t *starts, const uint32_t *lens, size_t count,
                     const double *pwm, size_t L, double del, double ext, int threads, double *f,
                     uint64_t *coords, uint64_t *aln_len, uint32_t *numbered, uint8_t *q_aln,
                     size_t slot, int32_t *status);

#ifdef __cplusplus
}
#endif
#endif
