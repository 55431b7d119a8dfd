// ref_shim16.cc -- the UNMODIFIED reference instantiated with two-byte position deltas:
// sblas::SparseMatrix<uint16, uint8, float> (code paths exist in the reference, sparse-matrix.cc:24,27, but its only
// instantiation is <uint8, uint8, float>, :315).
//
// TEST INFRASTRUCTURE, like ref_shim.cc.  The reference keeps its template bodies in the .cc files, so this translation
// unit includes src/sparse/kernel.cc and src/sparse/sparse-matrix.cc from where they lie under /root/reference (through
// the include path; nothing is copied into the repo) and adds the two explicit instantiations the uint16 class needs.
// Built by oracle/build_ref.sh into its own library, oracle/_ref/libsparsematrix_ref16.so.
#include <cstdint>
#include <cstring>
#include <vector>
#define private public
#include "sparse-matrix.h"
#undef private
#include "kernel.cc"
#include "sparse-matrix.cc"

template void sblas_kernel_operation_trans_ex<uint16_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(
    int m, int n, int k, float *a, int lda, float *c, int ldc, float alpha, uint16_t *ppos, uint8_t *pval, int pos_len,
    float *val_table, int valid_table_size);
namespace sblas {
template class SparseMatrix<uint16, uint8, float>;
}

#define REF_API extern "C" __attribute__((visibility("default")))

typedef sblas::SparseMatrix<uint16, uint8, float> Ref16;

REF_API void *sref16_create() { Ref16 *p = new Ref16(); p->Destroy(); return p; }
REF_API void sref16_free(void *h) { delete static_cast<Ref16 *>(h); }
REF_API void sref16_copyform(void *h, const uint8_t *codes, int32_t rows, int32_t cols, int32_t stride, const float *vals,
                             int32_t T, int32_t trans) {
    static_cast<Ref16 *>(h)->CopyForm(codes, rows, cols, stride, vals, T, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
}
REF_API void sref16_copyto(void *h, float *dense, int32_t stride, int32_t trans) {
    static_cast<Ref16 *>(h)->CopyTo(dense, stride, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
}
REF_API void sref16_addmatmat(void *h, float *a, int32_t m, int32_t lda, float *c, int32_t ldc, float alpha, float beta) {
    static_cast<Ref16 *>(h)->AddMatMat(a, m, lda, c, ldc, alpha, beta);
}
REF_API int32_t sref16_rows(void *h) { return static_cast<Ref16 *>(h)->NumRows(); }
REF_API int32_t sref16_cols(void *h) { return static_cast<Ref16 *>(h)->NumCols(); }
REF_API int64_t sref16_entries(void *h) { return (int64_t)static_cast<Ref16 *>(h)->pos_index_.size(); }
REF_API int32_t sref16_nblocks(void *h) { return (int32_t)static_cast<Ref16 *>(h)->block_bounds_.size(); }
REF_API int32_t sref16_table_len(void *h) { return (int32_t)static_cast<Ref16 *>(h)->val_table_.size(); }
REF_API void sref16_export(void *h, uint16_t *pos, uint8_t *code, float *table, int32_t *block_bounds,
                           int32_t *block_index_bounds) {
    Ref16 *p = static_cast<Ref16 *>(h);
    if (!p->pos_index_.empty()) {
        memcpy(pos, p->pos_index_.data(), p->pos_index_.size() * sizeof(uint16_t));
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
