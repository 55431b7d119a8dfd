// host_capi.cc -- extern "C" handle API over sblas::SparseMatrix (include/sblas_b200_host.h).
// This translation unit is the class's only friend-less client that needs the compressed
// members, so it reads them through a layout-identical view (the members are private, as in
// the reference, sparse-matrix.h:45-52).
#include <stdio.h>

#include <new>

#include "sblas_b200_host.h"
#include "sparse-matrix.h"

typedef sblas::SparseMatrix<uint8, uint8, float> HostMatrix;

namespace {
// Same member sequence as HostMatrix; checked by size below and by the structure parity tests.
struct MemberView {
    std::vector<uint8_t> pos_index;
    std::vector<uint8_t> val_index;
    std::vector<float> val_table;
    std::vector<std::pair<int32, int32> > block_bounds;
    std::vector<std::pair<int32, int32> > block_index_bounds;
    int32 rows, cols;
};
static_assert(sizeof(MemberView) == sizeof(HostMatrix), "SparseMatrix layout changed");
inline const MemberView &view(const sblas_host_matrix *h) { return *reinterpret_cast<const MemberView *>(h); }
inline HostMatrix *obj(sblas_host_matrix *h) { return reinterpret_cast<HostMatrix *>(h); }
inline const HostMatrix *obj(const sblas_host_matrix *h) { return reinterpret_cast<const HostMatrix *>(h); }
}  // namespace

extern "C" {

sblas_host_matrix *sblas_host_create(void) { return reinterpret_cast<sblas_host_matrix *>(new (std::nothrow) HostMatrix()); }
void sblas_host_free(sblas_host_matrix *h) { delete obj(h); }
void sblas_host_destroy(sblas_host_matrix *h) { if (h) obj(h)->Destroy(); }

int sblas_host_copyform(sblas_host_matrix *h, const uint8_t *density_matrix, int32_t rows, int32_t cols, int32_t stride,
                        const float *vals, int32_t val_table_size, int32_t trans) {
    if (!h || val_table_size < 0 || val_table_size > 255) return SBLAS_B200_EINVAL;   // the class aborts like the reference asserts
    obj(h)->CopyForm(density_matrix, rows, cols, stride, vals, val_table_size, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
    return SBLAS_B200_OK;
}
void sblas_host_copyto(sblas_host_matrix *h, float *density_matrix, int32_t stride, int32_t trans) {
    obj(h)->CopyTo(density_matrix, stride, trans ? sblas::SblasTrans : sblas::SblasNoTrans);
}
int sblas_host_addmatmat(sblas_host_matrix *h, const float *a, int32_t m, int32_t lda, float *c, int32_t ldc, float alpha,
                         float beta) {
    if (!h) return SBLAS_B200_EINVAL;
    if (obj(h)->NumCols() == 0 || m <= 0) return SBLAS_B200_OK;
    sblas_b200_matrix *dev = obj(h)->DeviceMatrix();
    if (!dev) return SBLAS_B200_ECUDA;
    return sblas_b200_addmatmat_host(dev, a, m, lda, c, ldc, alpha, beta);
}
int32_t sblas_host_num_rows(const sblas_host_matrix *h) { return obj(h)->NumRows(); }
int32_t sblas_host_num_cols(const sblas_host_matrix *h) { return obj(h)->NumCols(); }
int32_t sblas_host_equal(sblas_host_matrix *a, sblas_host_matrix *b) { return (*obj(a) == *obj(b)) ? 1 : 0; }
int32_t sblas_host_selftest(void) { HostMatrix t; return t.SelfTest() ? 1 : 0; }

int64_t sblas_host_entries(const sblas_host_matrix *h) { return (int64_t)view(h).pos_index.size(); }
int32_t sblas_host_nblocks(const sblas_host_matrix *h) { return (int32_t)view(h).block_bounds.size(); }
int32_t sblas_host_table_size(const sblas_host_matrix *h) {
    return view(h).val_table.empty() ? 0 : (int32_t)view(h).val_table.size() - 1;
}
void sblas_host_members(const sblas_host_matrix *h, const uint8_t **pos_index, const uint8_t **val_index,
                        const float **val_table, const int32_t **block_bounds, const int32_t **block_index_bounds) {
    const MemberView &v = view(h);
    if (pos_index) *pos_index = v.pos_index.data();
    if (val_index) *val_index = v.val_index.data();
    if (val_table) *val_table = v.val_table.data();
    if (block_bounds) *block_bounds = v.block_bounds.empty() ? nullptr : &v.block_bounds[0].first;
    if (block_index_bounds) *block_index_bounds = v.block_index_bounds.empty() ? nullptr : &v.block_index_bounds[0].first;
}
sblas_b200_matrix *sblas_host_device_matrix(sblas_host_matrix *h) { return h ? obj(h)->DeviceMatrix() : nullptr; }

}  // extern "C"
