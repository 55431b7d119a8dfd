// kernel.cc -- the reference's free kernel functions bound to the CUDA library (see kernel.h).
#include "kernel.h"

#include <stdio.h>

#include "sblas_b200.h"

#define SBLAS_VISIBLE __attribute__((visibility("default")))

static void report(const char *what, int rc) {
    if (rc != SBLAS_B200_OK) printf("%s failed (%d): %s\n", what, rc, sblas_b200_last_error());
}

template <>
SBLAS_VISIBLE void sblas_beta_operation_kernel<float>(float *c, int m, int n, int ldc, float beta) {
    report("sblas_beta_operation_kernel", sblas_b200_scale_host(c, m, n, ldc, beta));
}

template <>
SBLAS_VISIBLE void sblas_trans_kernel<float>(float *a, int m, int n, int lda, float *sa, int ldsa) {
    report("sblas_trans_kernel", sblas_b200_transpose_host(a, m, n, lda, sa, ldsa));
}

#define SBLAS_PANEL(NAME, TRANS)                                                                                        \
    template <>                                                                                                         \
    SBLAS_VISIBLE void NAME<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(                                            \
        int m, int n, int k, float *a, int lda, float *c, int ldc, float alpha, uint8_t *ppos, uint8_t *pval, int pos_len, \
        float *val_table, int valid_table_size) {                                                                       \
        report(#NAME, sblas_b200_panel_product_host(TRANS, m, n, k, a, lda, c, ldc, alpha, ppos, pval, pos_len, val_table,  \
                                                    valid_table_size));                                                 \
    }

SBLAS_PANEL(sblas_kernel_operation, 0)
SBLAS_PANEL(sblas_kernel_operation_naive, 0)
SBLAS_PANEL(sblas_kernel_operation_trans, 1)
SBLAS_PANEL(sblas_kernel_operation_trans_ex, 1)
