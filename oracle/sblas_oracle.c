/*
 * sblas_oracle.c -- CPU restatement of the NeverLEX/sparsematrix sparse x dense
 * product path, in plain C.
 *
 * THIS FILE IS TEST INFRASTRUCTURE.  It is the checker that the parity tests
 * (tests/), __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs compare the CUDA path against.  Nothing under
 * sparsematrix_b200/ may include, link or call it.
 *
 * Parity status: PINNED.  Every function below is checked (tests/test_oracle_*.py)
 *   - against the reference's own known-answer tests (SelfTest KAT-1 / KAT-2,
 *     sparse-matrix.cc:211-246) and its 1023x511 round trip (:247-311),
 *   - bit-for-bit against the unmodified reference library compiled from
 *     /root/reference into oracle/_ref/ (oracle/build_ref.sh), and
 *   - against the committed golden dumps in tests/golden/ that were produced
 *     by that reference build (tests/golden/make_golden.py).
 *
 * Build: oracle/Makefile  (gcc -O2 -ffp-contract=off: the reference's generic
 * build rounds the product and the sum separately -- no FMA -- and so must we).
 *
 * Notation (reference): C[m x n] = beta*C + alpha * A[m x k] * S[k x n],
 *   A, C dense row-major, S sparse in the compressed panel format:
 *   column panels of 256, per stored entry one byte of position delta and one
 *   byte of codebook index; see SURVEY.md appendix A.
 */
#ifdef _OPENMP
#include <omp.h>
#endif
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#define SBO_PANEL_SHIFT 8                 /* kernel.h:26  SBLAS_BLOCK_COL_SHIFT */
#define SBO_PANEL_WIDTH (1 << SBO_PANEL_SHIFT)
#define SBO_PANEL_MASK  (SBO_PANEL_WIDTH - 1)
#define SBO_MAX_DELTA   255               /* sparse-matrix.cc:24  zero_pad_interval for uint8 */

#define SBO_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------- *
 * Format build.  Follows sparse-matrix.cc:21-99 (CopyForm).
 *
 * S(r, j) for the two orientations:
 *   trans == 0 : S is rows x cols,  S(r,j) = codes[r*stride + j]      (:32-64)
 *   trans == 1 : S is cols x rows,  S(r,j) = codes[j*stride + r]      (:65-98)
 * A code is a stored non-zero iff code < T (:44, :77).  Panels walk the
 * columns of S in steps of 256 and span every row (block_row_shift == 0, :27).
 *
 * Two-call protocol: call with pos == NULL to obtain the sizes, then again
 * with buffers.  Returns the number of stream entries (fillers included);
 * *nblocks_out = number of non-empty panels.
 * ------------------------------------------------------------------------- */
/* position deltas are PosIndex_t: one byte in the reference's only instantiation (sparse-matrix.cc:315), two bytes in
 * SparseMatrix<uint16, uint8, float> -- the same code paths with zero_pad_interval = 65535 (:24); pb = bytes per delta */
static inline uint32_t rd_pos(const void *pos, int pb, int64_t i)
{
    return pb == 1 ? (uint32_t)((const uint8_t *)pos)[i] : (uint32_t)((const uint16_t *)pos)[i];
}
static inline void wr_pos(void *pos, int pb, int64_t i, uint32_t v)
{
    if (pb == 1) ((uint8_t *)pos)[i] = (uint8_t)v; else ((uint16_t *)pos)[i] = (uint16_t)v;
}

static int64_t copyform_impl(const uint8_t *codes, int32_t rows, int32_t cols, int32_t stride,
                             const float *vals, int32_t T, int32_t trans, int pb,
                             void *pos, uint8_t *code, float *table /* T+1 */,
                             int32_t *block_bounds /* 2*nblocks: (row_off, col_off) */,
                             int32_t *block_index_bounds /* 2*nblocks: [start, end) */,
                             int32_t *nblocks_out, int32_t *s_rows_out, int32_t *s_cols_out)
{
    int64_t n_entries = 0;
    int32_t n_blocks = 0;
    const int32_t max_delta = pb == 1 ? SBO_MAX_DELTA : 65535;   /* :24 zero_pad_interval */
    if (T < 0 || T > SBO_MAX_DELTA) return -1;          /* :25 assert (codes are one byte in both instantiations) */
    if (T == 0) {                                       /* :26 empty object, rows_/cols_ stay 0 */
        if (nblocks_out) *nblocks_out = 0;
        if (s_rows_out) *s_rows_out = 0;
        if (s_cols_out) *s_cols_out = 0;
        return 0;
    }
    const int32_t s_rows = trans ? cols : rows;         /* :63-64, :96-97 */
    const int32_t s_cols = trans ? rows : cols;
    if (table) {                                        /* :29-31 */
        memcpy(table, vals, (size_t)T * sizeof(float));
        table[T] = 0.0f;
    }
    for (int32_t j0 = 0; j0 < s_cols; j0 += SBO_PANEL_WIDTH) {
        const int32_t width = (s_cols - j0 >= SBO_PANEL_WIDTH) ? SBO_PANEL_WIDTH : s_cols - j0;
        const int64_t left = n_entries;
        int32_t prev = 0;                               /* :36 prev_index */
        for (int32_t r = 0; r < s_rows; ++r) {
            for (int32_t jj = 0; jj < width; ++jj) {
                const uint8_t cd = trans ? codes[(int64_t)(j0 + jj) * stride + r]
                                         : codes[(int64_t)r * stride + (j0 + jj)];
                if (cd >= T) continue;                  /* :44 */
                const int32_t lin = r * SBO_PANEL_WIDTH + jj;   /* :41 stride is 256 even in a narrow panel */
                int32_t d = lin - prev;
                while (d > max_delta) {                 /* :46-51 filler */
                    if (pos) { wr_pos(pos, pb, n_entries, (uint32_t)max_delta); code[n_entries] = (uint8_t)T; }
                    ++n_entries;
                    d -= max_delta;
                }
                if (pos) { wr_pos(pos, pb, n_entries, (uint32_t)d); code[n_entries] = cd; }
                ++n_entries;
                prev = lin;
            }
        }
        if (n_entries != left) {                        /* :57-60 only non-empty panels are recorded */
            if (block_bounds) { block_bounds[2 * n_blocks] = 0; block_bounds[2 * n_blocks + 1] = j0; }
            if (block_index_bounds) {
                block_index_bounds[2 * n_blocks] = (int32_t)left;
                block_index_bounds[2 * n_blocks + 1] = (int32_t)n_entries;
            }
            ++n_blocks;
        }
    }
    if (nblocks_out) *nblocks_out = n_blocks;
    if (s_rows_out) *s_rows_out = s_rows;
    if (s_cols_out) *s_cols_out = s_cols;
    return n_entries;
}

SBO_API int64_t sbo_copyform(const uint8_t *codes, int32_t rows, int32_t cols, int32_t stride,
                             const float *vals, int32_t T, int32_t trans,
                             uint8_t *pos, uint8_t *code, float *table, int32_t *block_bounds,
                             int32_t *block_index_bounds, int32_t *nblocks_out, int32_t *s_rows_out, int32_t *s_cols_out)
{
    return copyform_impl(codes, rows, cols, stride, vals, T, trans, 1, pos, code, table, block_bounds, block_index_bounds,
                         nblocks_out, s_rows_out, s_cols_out);
}
SBO_API int64_t sbo_copyform16(const uint8_t *codes, int32_t rows, int32_t cols, int32_t stride,
                               const float *vals, int32_t T, int32_t trans,
                               uint16_t *pos, uint8_t *code, float *table, int32_t *block_bounds,
                               int32_t *block_index_bounds, int32_t *nblocks_out, int32_t *s_rows_out, int32_t *s_cols_out)
{
    return copyform_impl(codes, rows, cols, stride, vals, T, trans, 2, pos, code, table, block_bounds, block_index_bounds,
                         nblocks_out, s_rows_out, s_cols_out);
}

/* ------------------------------------------------------------------------- *
 * Decode to dense.  Follows sparse-matrix.cc:101-137 (CopyTo).
 *   trans == 0 : dense is s_rows x stride, dense[r*stride + j] = S(r,j)
 *   trans == 1 : dense is s_cols x stride, dense[j*stride + r] = S(r,j)
 * Note the reference clears rows_*stride (resp. cols_*stride) floats first.
 * ------------------------------------------------------------------------- */
static void copyto_impl(const void *pos, int pb, const uint8_t *code, const float *table, int32_t T,
                        const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                        int32_t s_rows, int32_t s_cols, float *dense, int32_t stride, int32_t trans)
{
    const int64_t lead = trans ? s_cols : s_rows;
    memset(dense, 0, (size_t)lead * stride * sizeof(float));            /* :109, :123 */
    for (int32_t b = 0; b < nblocks; ++b) {
        const int32_t row_off = block_bounds[2 * b], col_off = block_bounds[2 * b + 1];
        int32_t lin = 0;
        for (int32_t e = block_index_bounds[2 * b]; e < block_index_bounds[2 * b + 1]; ++e) {
            lin += (int32_t)rd_pos(pos, pb, e);                         /* :116 */
            if (code[e] == T) continue;                                 /* :117 */
            const int32_t r = row_off + (lin >> SBO_PANEL_SHIFT), j = col_off + (lin & SBO_PANEL_MASK);
            if (trans) dense[(int64_t)j * stride + r] = table[code[e]]; /* :133 */
            else       dense[(int64_t)r * stride + j] = table[code[e]]; /* :119 */
        }
    }
}

SBO_API void sbo_copyto(const uint8_t *pos, const uint8_t *code, const float *table, int32_t T,
                        const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                        int32_t s_rows, int32_t s_cols, float *dense, int32_t stride, int32_t trans)
{
    copyto_impl(pos, 1, code, table, T, block_bounds, block_index_bounds, nblocks, s_rows, s_cols, dense, stride, trans);
}
SBO_API void sbo_copyto16(const uint16_t *pos, const uint8_t *code, const float *table, int32_t T,
                          const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                          int32_t s_rows, int32_t s_cols, float *dense, int32_t stride, int32_t trans)
{
    copyto_impl(pos, 2, code, table, T, block_bounds, block_index_bounds, nblocks, s_rows, s_cols, dense, stride, trans);
}

/* ------------------------------------------------------------------------- *
 * The product, fp32, in the reference's arithmetic order.
 *
 * AddMatMat (sparse-matrix.cc:139-194) does: if (beta != 1) C *= beta
 * (kernel.cc:10-29); if (alpha != 0) for each recorded panel: transpose in,
 * sblas_kernel_operation_trans_ex (kernel.cc:771-800), transpose out.  The
 * transposes (kernel.cc:31-187) move values without arithmetic, and the row
 * batching of trans_ex only regroups the same updates, so for every output
 * element C[i][j] the value is
 *     acc = (beta != 1) ? fl(C[i][j]*beta) : C[i][j]
 *     for stored entries of column j in ascending r:
 *         acc = fl(acc + fl(A[i][r] * fl(table[code]*alpha)))      (kernel.cc:791-794, :568-582)
 * which is what this loop nest computes, directly on the untransposed arrays
 * (the same nest as the reference's own sblas_kernel_operation_naive,
 * kernel.cc:322-338).  Entries with code >= T are fillers (kernel.cc:782).
 * ------------------------------------------------------------------------- */
SBO_API void sbo_addmatmat(const uint8_t *pos, const uint8_t *code, const float *table, int32_t T,
                           const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                           int32_t s_rows /* k */, int32_t s_cols /* n */,
                           const float *a, int32_t m, int32_t lda, float *c, int32_t ldc,
                           float alpha, float beta)
{
    (void)s_rows;
    if (beta != 1.0) {                                  /* sparse-matrix.cc:149 (compares against a double literal) */
        for (int32_t i = 0; i < m; ++i)
            for (int32_t j = 0; j < s_cols; ++j) c[(int64_t)i * ldc + j] *= beta;
    }
    if (alpha == 0.0) return;                           /* :152 */
    for (int32_t b = 0; b < nblocks; ++b) {
        const int32_t row_off = block_bounds[2 * b], col_off = block_bounds[2 * b + 1];
        int32_t lin = 0;
        for (int32_t e = block_index_bounds[2 * b]; e < block_index_bounds[2 * b + 1]; ++e) {
            lin += pos[e];
            if (code[e] >= T) continue;
            const int32_t r = row_off + (lin >> SBO_PANEL_SHIFT), j = col_off + (lin & SBO_PANEL_MASK);
            const float w = table[code[e]] * alpha;     /* alpha is folded into the weight first */
            for (int32_t i = 0; i < m; ++i)
                c[(int64_t)i * ldc + j] += a[(int64_t)i * lda + r] * w;
        }
    }
}

/* fp64 statement of the same product plus the backward-error denominator
 *   y64[i][j]   = beta*C + alpha * sum_r A[i][r]*table[code]
 *   bound[i][j] = |beta*C| + |alpha| * sum_r |A[i][r]|*|table[code]|
 * (err = max |y - y64| / bound <= 1e-5).  Outputs are m x n,
 * row-major, leading dimension n. */
SBO_API void sbo_addmatmat_f64(const uint8_t *pos, const uint8_t *code, const float *table, int32_t T,
                               const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                               int32_t s_rows, int32_t s_cols,
                               const float *a, int32_t m, int32_t lda, const float *c, int32_t ldc,
                               float alpha, float beta, double *y64, double *bound)
{
    (void)s_rows;
    for (int32_t i = 0; i < m; ++i)
        for (int32_t j = 0; j < s_cols; ++j) {
            const double v = (double)beta * (double)c[(int64_t)i * ldc + j];
            y64[(int64_t)i * s_cols + j] = v;
            if (bound) bound[(int64_t)i * s_cols + j] = fabs(v);
        }
    for (int32_t b = 0; b < nblocks; ++b) {
        const int32_t row_off = block_bounds[2 * b], col_off = block_bounds[2 * b + 1];
        int32_t lin = 0;
        for (int32_t e = block_index_bounds[2 * b]; e < block_index_bounds[2 * b + 1]; ++e) {
            lin += pos[e];
            if (code[e] >= T) continue;
            const int32_t r = row_off + (lin >> SBO_PANEL_SHIFT), j = col_off + (lin & SBO_PANEL_MASK);
            const double w = (double)table[code[e]] * (double)alpha;
            for (int32_t i = 0; i < m; ++i) {
                const double p = (double)a[(int64_t)i * lda + r] * w;
                y64[(int64_t)i * s_cols + j] += p;
                if (bound) bound[(int64_t)i * s_cols + j] += fabs(p);
            }
        }
    }
}

/* ------------------------------------------------------------------------- *
 * Structure oracle for the device upload: the compressed stream rewritten as
 * CSR of W = S^T (W row j = column j of S; inside a row the entries keep the
 * reference's accumulation order, ascending r).  Derived from the decode loop
 * kernel.cc:780-796 / sparse-matrix.cc:110-120.  row_ptr has s_cols+1 entries.
 * Returns nnz (stored non-zeros, fillers dropped); *fillers_out counts fillers.
 * Call with col == NULL to size.
 * ------------------------------------------------------------------------- */
static int64_t stream_to_csr_impl(const void *pos, int pb, const uint8_t *code, int32_t T,
                                  const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                                  int32_t s_rows, int32_t s_cols,
                                  int32_t *row_ptr, int32_t *col, uint8_t *csr_code, int64_t *fillers_out)
{
    (void)s_rows;
    int64_t nnz = 0, fillers = 0;
    int32_t *cnt = (int32_t *)calloc((size_t)s_cols + 1, sizeof(int32_t));
    if (!cnt) return -1;
    for (int32_t b = 0; b < nblocks; ++b) {
        const int32_t col_off = block_bounds[2 * b + 1];
        int32_t lin = 0;
        for (int32_t e = block_index_bounds[2 * b]; e < block_index_bounds[2 * b + 1]; ++e) {
            lin += (int32_t)rd_pos(pos, pb, e);
            if (code[e] >= T) { ++fillers; continue; }
            ++cnt[col_off + (lin & SBO_PANEL_MASK)];
            ++nnz;
        }
    }
    if (fillers_out) *fillers_out = fillers;
    if (row_ptr) {
        int32_t run = 0;
        for (int32_t j = 0; j < s_cols; ++j) { row_ptr[j] = run; run += cnt[j]; }
        row_ptr[s_cols] = run;
    }
    if (col && row_ptr) {
        int32_t *cur = (int32_t *)malloc(((size_t)s_cols + 1) * sizeof(int32_t));
        if (!cur) { free(cnt); return -1; }
        memcpy(cur, row_ptr, ((size_t)s_cols + 1) * sizeof(int32_t));
        for (int32_t b = 0; b < nblocks; ++b) {
            const int32_t row_off = block_bounds[2 * b], col_off = block_bounds[2 * b + 1];
            int32_t lin = 0;
            for (int32_t e = block_index_bounds[2 * b]; e < block_index_bounds[2 * b + 1]; ++e) {
                lin += (int32_t)rd_pos(pos, pb, e);
                if (code[e] >= T) continue;
                const int32_t j = col_off + (lin & SBO_PANEL_MASK);
                col[cur[j]] = row_off + (lin >> SBO_PANEL_SHIFT);
                csr_code[cur[j]] = code[e];
                ++cur[j];
            }
        }
        free(cur);
    }
    free(cnt);
    return nnz;
}

SBO_API int64_t sbo_stream_to_csr(const uint8_t *pos, const uint8_t *code, int32_t T,
                                  const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                                  int32_t s_rows, int32_t s_cols,
                                  int32_t *row_ptr, int32_t *col, uint8_t *csr_code, int64_t *fillers_out)
{
    return stream_to_csr_impl(pos, 1, code, T, block_bounds, block_index_bounds, nblocks, s_rows, s_cols, row_ptr, col, csr_code, fillers_out);
}
SBO_API int64_t sbo_stream_to_csr16(const uint16_t *pos, const uint8_t *code, int32_t T,
                                    const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                                    int32_t s_rows, int32_t s_cols,
                                    int32_t *row_ptr, int32_t *col, uint8_t *csr_code, int64_t *fillers_out)
{
    return stream_to_csr_impl(pos, 2, code, T, block_bounds, block_index_bounds, nblocks, s_rows, s_cols, row_ptr, col, csr_code, fillers_out);
}

/* ------------------------------------------------------------------------- *
 * The same product stated on CSR(W) input -- needed for shapes the panel
 * format cannot hold (BASELINE.json config 4: 2^23 rows) and
 * for the direct-CSR upload entry point.  Arithmetic is the reference's
 * (kernel.cc:791-794, :568-582): weight = fl(table[code]*alpha), then for
 * each output element the updates are applied in ascending r (= CSR order
 * inside a row), product and sum rounded separately.  Validated against
 * sbo_addmatmat / the reference on every shape both can run.
 * ------------------------------------------------------------------------- */
SBO_API void sbo_addmatmat_csr(const int32_t *row_ptr, const int32_t *col, const uint8_t *csr_code,
                               const float *table, int32_t T, int32_t n /* W rows */, int32_t k,
                               const float *a, int32_t m, int32_t lda, float *c, int32_t ldc,
                               float alpha, float beta)
{
    (void)k; (void)T;
    if (beta != 1.0) {
        for (int32_t i = 0; i < m; ++i)
            for (int32_t j = 0; j < n; ++j) c[(int64_t)i * ldc + j] *= beta;
    }
    if (alpha == 0.0) return;
    float w_alpha[256];
    for (int t = 0; t < 256; ++t) w_alpha[t] = (t < T) ? table[t] * alpha : 0.0f;
    if (m == 1) {
        for (int32_t j = 0; j < n; ++j) {
            float acc = c[j];
            for (int32_t e = row_ptr[j]; e < row_ptr[j + 1]; ++e) acc += a[col[e]] * w_alpha[csr_code[e]];
            c[j] = acc;
        }
        return;
    }
    for (int32_t j = 0; j < n; ++j)
        for (int32_t e = row_ptr[j]; e < row_ptr[j + 1]; ++e) {
            const float w = w_alpha[csr_code[e]];
            const int32_t r = col[e];
            for (int32_t i = 0; i < m; ++i) c[(int64_t)i * ldc + j] += a[(int64_t)i * lda + r] * w;
        }
}

/* The m == 1 case with the rows spread over all host cores (OpenMP).  NOT the reference -- the reference is
 * single-threaded -- but the fairest CPU number for the 2^23-row power-law case, which the reference's own
 * format cannot hold.  Every row is summed exactly as above, so the bits are the same. */
SBO_API int32_t sbo_addmatmat_csr_omp(const int32_t *row_ptr, const int32_t *col, const uint8_t *csr_code,
                                      const float *table, int32_t T, int32_t n, const float *x, float *y,
                                      float alpha, float beta)
{
    float w_alpha[256];
    for (int t = 0; t < 256; ++t) w_alpha[t] = (t < T) ? table[t] * alpha : 0.0f;
    int32_t threads = 1;
#ifdef _OPENMP
    threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4096)
#endif
    for (int32_t j = 0; j < n; ++j) {
        float acc = y[j];
        if (beta != 1.0) acc *= beta;
        if (alpha != 0.0)
            for (int32_t e = row_ptr[j]; e < row_ptr[j + 1]; ++e) acc += x[col[e]] * w_alpha[csr_code[e]];
        y[j] = acc;
    }
    return threads;
}

SBO_API void sbo_addmatmat_csr_f64(const int32_t *row_ptr, const int32_t *col, const uint8_t *csr_code,
                                   const float *table, int32_t T, int32_t n, int32_t k,
                                   const float *a, int32_t m, int32_t lda, const float *c, int32_t ldc,
                                   float alpha, float beta, double *y64, double *bound)
{
    (void)k;
    for (int32_t i = 0; i < m; ++i)
        for (int32_t j = 0; j < n; ++j) {
            double acc = (double)beta * (double)c[(int64_t)i * ldc + j], bd = fabs(acc);
            for (int32_t e = row_ptr[j]; e < row_ptr[j + 1]; ++e) {
                if (csr_code[e] >= T) continue;
                const double p = (double)a[(int64_t)i * lda + col[e]] * ((double)table[csr_code[e]] * (double)alpha);
                acc += p; bd += fabs(p);
            }
            y64[(int64_t)i * n + j] = acc;
            if (bound) bound[(int64_t)i * n + j] = bd;
        }
}

/* Out-of-place transpose, the function sblas_trans_kernel computes
 * (kernel.cc:31-187): sa[j*ldsa + i] = a[i*lda + j], i < m, j < n.
 * kernel_test.cc:5-12 states the same thing as its own check. */
SBO_API void sbo_transpose(const float *a, int32_t m, int32_t n, int32_t lda, float *sa, int32_t ldsa)
{
    for (int32_t i = 0; i < m; ++i)
        for (int32_t j = 0; j < n; ++j) sa[(int64_t)j * ldsa + i] = a[(int64_t)i * lda + j];
}

/* Error measures used by the parity tests:
 *   out[0] = max |y - y64| / bound      (backward error; the <= 1e-5 contract)
 *   out[1] = max |y - y64| / max |y64|  (normwise)
 *   out[2] = max |y - y64| / |y64|      (elementwise, informational)
 * y is m x n with leading dimension ldy; y64/bound are m x n dense. */
SBO_API void sbo_error_measures(const float *y, int32_t m, int32_t n, int32_t ldy,
                                const double *y64, const double *bound, double *out)
{
    double e_back = 0, e_abs = 0, y_max = 0, e_elem = 0;
    for (int32_t i = 0; i < m; ++i)
        for (int32_t j = 0; j < n; ++j) {
            const double ref = y64[(int64_t)i * n + j];
            const double d = fabs((double)y[(int64_t)i * ldy + j] - ref);
            const double bd = bound[(int64_t)i * n + j];
            if (bd > 0 && d / bd > e_back) e_back = d / bd;
            if (bd == 0 && d != 0) e_back = INFINITY;
            if (d > e_abs) e_abs = d;
            if (fabs(ref) > y_max) y_max = fabs(ref);
            if (ref != 0 && d / fabs(ref) > e_elem) e_elem = d / fabs(ref);
        }
    out[0] = e_back;
    out[1] = y_max > 0 ? e_abs / y_max : e_abs;
    out[2] = e_elem;
}
