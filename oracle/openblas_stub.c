/* openblas_stub.c -- the two non-standard entry points of the reference's PATCHED OpenBLAS 0.3.4.dev
 * (openblas_arm64_plus/include/cblas.h:259-262), which its blas_test.cc:21-29 calls for the third row of
 * its table.  The patched library ships only as ARM binaries, so the unmodified blas_test.cc is linked
 * against a stock x86-64 OpenBLAS plus these two functions: precopy = plain copy,
 * mul = cblas_sgemm.  Test infrastructure only; the sparse path never touches OpenBLAS. */
#include <string.h>

enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113, CblasConjNoTrans = 114 };

void cblas_sgemm(enum CBLAS_ORDER, enum CBLAS_TRANSPOSE, enum CBLAS_TRANSPOSE, int, int, int, float, const float *, int,
                 const float *, int, float, float *, int);

void cblas_sgemm_precopy(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta, enum CBLAS_TRANSPOSE tb, int m, int k,
                         const float *a, int lda, float *c) {
    (void)order; (void)ta; (void)tb;
    for (int i = 0; i < m; ++i) memcpy(c + (size_t)i * k, a + (size_t)i * lda, (size_t)k * sizeof(float));
}

void cblas_sgemm_mul(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta, enum CBLAS_TRANSPOSE tb, int m, int n, int k,
                     float alpha, const float *a, int lda, const float *b, int ldb, float beta, float *c, int ldc) {
    cblas_sgemm(order, ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc);
}
