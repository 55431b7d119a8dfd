// kernel.h -- the reference's free kernel functions (reference src/sparse/kernel.h:42-62), same template
// names, same signatures, same explicit instantiations (kernel.cc:802-811 there) -- so code written
// against the reference header links against this build -- backed by the CUDA library
// (include/sblas_b200.h: sblas_b200_scale_host, _transpose_host, _panel_product_host).
// Error convention as in the reference: void, a message on stdout and return.
#ifndef SBLAS_B200_KERNEL_H_
#define SBLAS_B200_KERNEL_H_

#include <stdint.h>

#ifndef SBLAS_BLOCK_ROW_SHIFT
#define SBLAS_BLOCK_ROW_SHIFT 0
#endif
#ifndef SBLAS_BLOCK_COL_SHIFT
#define SBLAS_BLOCK_COL_SHIFT 8
#endif

// c[m x n, ldc] *= beta                                              (kernel.cc:10-29)
template <typename type_t>
void sblas_beta_operation_kernel(type_t *c, int m, int n, int ldc, type_t beta);

// sa[j * ldsa + i] = a[i * lda + j], i < m, j < n                    (kernel.cc:31-187)
template <typename type_t>
void sblas_trans_kernel(type_t *a, int m, int n, int lda, type_t *sa, int ldsa);

// C += alpha * A * P for one panel P (k x n) given as its byte streams; a: m x k, c: m x n, row-major   (kernel.cc:213-369)
template <typename PosIndex_t, typename ValIndex_t, typename Value_t, const int block_col_shift>
void sblas_kernel_operation(int m, int n, int k, Value_t *a, int lda, Value_t *c, int ldc, Value_t alpha, PosIndex_t *ppos,
                            ValIndex_t *pval, int pos_len, Value_t *val_table, int valid_table_size);
template <typename PosIndex_t, typename ValIndex_t, typename Value_t, const int block_col_shift>
void sblas_kernel_operation_naive(int m, int n, int k, Value_t *a, int lda, Value_t *c, int ldc, Value_t alpha,
                                  PosIndex_t *ppos, ValIndex_t *pval, int pos_len, Value_t *val_table, int valid_table_size);

// the same on transposed operands: a is k x lda, c is n x ldc, the m values contiguous   (kernel.cc:612-800)
template <typename PosIndex_t, typename ValIndex_t, typename Value_t, const int block_col_shift>
void sblas_kernel_operation_trans(int m, int n, int k, Value_t *a, int lda, Value_t *c, int ldc, Value_t alpha,
                                  PosIndex_t *ppos, ValIndex_t *pval, int pos_len, Value_t *val_table, int valid_table_size);
template <typename PosIndex_t, typename ValIndex_t, typename Value_t, const int block_col_shift>
void sblas_kernel_operation_trans_ex(int m, int n, int k, Value_t *a, int lda, Value_t *c, int ldc, Value_t alpha,
                                     PosIndex_t *ppos, ValIndex_t *pval, int pos_len, Value_t *val_table, int valid_table_size);

#endif
