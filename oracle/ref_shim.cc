// ref_shim.cc -- extern "C" handle API over the UNMODIFIED reference library.
//
// TEST INFRASTRUCTURE.  Compiled by oracle/build_ref.sh together with the
// reference's own src/sparse/kernel.cc and src/sparse/sparse-matrix.cc, read
// from where they lie under /root/reference (never copied into this repo),
// into oracle/_ref/libsparsematrix_ref.so.  It exists so that the Python
// tests and bench.py can (a) validate oracle/sblas_oracle.c bit-for-bit
// against the real thing, (b) generate tests/golden/, and (c) time the
// reference's CPU path (bench.py cpu_baseline kind "reference").
//
// The compressed members are private with no accessors (sparse-matrix.h:45-52),
// so this one translation unit re-labels `private` to read them.  The class
// layout is unchanged (GCC lays members out in declaration order whatever the
// access label), and the reference objects are compiled without the re-label.
#include <cstdint>
#include <cstring>
#include <vector>
#define private public
#include "sparse-matrix.h"
#undef private

#define REF_API extern "C" __attribute__((visibility("default")))

typedef sblas::SparseMatrix<uint8, uint8, float> RefMatrix;

REF_API void *sref_create() { RefMatrix *p = new RefMatrix(); p->Destroy(); return p; }
REF_API void sref_free(void *h) { delete static_cast<RefMatrix *>(h); }

REF_API void sref_copyform(void *h, const uint8_t *codes, int32_t rows, int32_t cols, int32_t stride,
                           const float *vals, int32_t T, int32_t trans) {
    static_cast<RefMatrix *>(h)->CopyForm(codes, rows, cols, stride, vals, T,
                                          trans ? sblas::SblasTrans : sblas::SblasNoTrans);
}
REF_API void sref_copyto(void *h, float *dense, int32_t stride, int32_t trans) {
    static_cast<RefMatrix *>(h)->CopyTo(dense, stride, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
}
REF_API void sref_addmatmat(void *h, float *a, int32_t m, int32_t lda, float *c, int32_t ldc,
                            float alpha, float beta) {
    static_cast<RefMatrix *>(h)->AddMatMat(a, m, lda, c, ldc, alpha, beta);
}
REF_API int32_t sref_rows(void *h) { return static_cast<RefMatrix *>(h)->NumRows(); }
REF_API int32_t sref_cols(void *h) { return static_cast<RefMatrix *>(h)->NumCols(); }
REF_API int64_t sref_entries(void *h) { return (int64_t)static_cast<RefMatrix *>(h)->pos_index_.size(); }
REF_API int32_t sref_nblocks(void *h) { return (int32_t)static_cast<RefMatrix *>(h)->block_bounds_.size(); }
REF_API int32_t sref_table_len(void *h) { return (int32_t)static_cast<RefMatrix *>(h)->val_table_.size(); }

// Copy the compressed members out: pos/code [entries], table [table_len],
// block_bounds / block_index_bounds as flat int32 pairs [2*nblocks].
REF_API void sref_export(void *h, uint8_t *pos, uint8_t *code, float *table,
                         int32_t *block_bounds, int32_t *block_index_bounds) {
    RefMatrix *p = static_cast<RefMatrix *>(h);
    if (!p->pos_index_.empty()) {
        memcpy(pos, p->pos_index_.data(), p->pos_index_.size());
        memcpy(code, p->val_index_.data(), p->val_index_.size());
    }
    if (!p->val_table_.empty()) memcpy(table, p->val_table_.data(), p->val_table_.size() * sizeof(float));
    for (size_t b = 0; b < p->block_bounds_.size(); ++b) {
        block_bounds[2 * b] = p->block_bounds_[b].first;
        block_bounds[2 * b + 1] = p->block_bounds_[b].second;
        block_index_bounds[2 * b] = p->block_index_bounds_[b].first;
        block_index_bounds[2 * b + 1] = p->block_index_bounds_[b].second;
    }
}

// The reference's built-in known-answer + round-trip test (sparse-matrix.cc:209-313).
REF_API int32_t sref_selftest() { RefMatrix t; t.Destroy(); return t.SelfTest() ? 1 : 0; }

// Free kernels (kernel.h:42-62), for cross-checks of the alternative formulations.
REF_API void sref_trans_kernel(float *a, int m, int n, int lda, float *sa, int ldsa) {
    sblas_trans_kernel<float>(a, m, n, lda, sa, ldsa);
}
REF_API void sref_beta_kernel(float *c, int m, int n, int ldc, float beta) {
    sblas_beta_operation_kernel<float>(c, m, n, ldc, beta);
}
// which: 0 = sblas_kernel_operation (8-row unrolled), 1 = _naive, 2 = _trans, 3 = _trans_ex
REF_API void sref_panel_kernel(int which, int m, int n, int k, float *a, int lda, float *c, int ldc, float alpha,
                               uint8_t *ppos, uint8_t *pval, int pos_len, float *val_table, int valid_table_size) {
    switch (which) {
    case 0: sblas_kernel_operation<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(m, n, k, a, lda, c, ldc, alpha, ppos, pval, pos_len, val_table, valid_table_size); break;
    case 1: sblas_kernel_operation_naive<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(m, n, k, a, lda, c, ldc, alpha, ppos, pval, pos_len, val_table, valid_table_size); break;
    case 2: sblas_kernel_operation_trans<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(m, n, k, a, lda, c, ldc, alpha, ppos, pval, pos_len, val_table, valid_table_size); break;
    default: sblas_kernel_operation_trans_ex<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(m, n, k, a, lda, c, ldc, alpha, ppos, pval, pos_len, val_table, valid_table_size); break;
    }
}
