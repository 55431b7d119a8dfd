// parity_test.cc -- GPU AddMatMat against an fp64 dense product, deterministic seeds, exit code != 0
// on failure.  The counterpart of the reference harness's check (src/test/blas_test.h:160-182,
// :240-245: AddMatMat vs cblas_sgemm on the CopyTo-expanded matrix, 10 % tolerance, message only)
// with the tight contract of this build: backward error <= 1e-5.
//
//   parity_test            the built-in sweep
//   parity_test m n k d    one case: m x n x k at density d
#include "harness.h"

using namespace harness;

static int run_case(int m, int n, int k, double density, float alpha, float beta, int pad_a, int pad_c, uint64_t seed) {
    const int T = 255;
    std::vector<float> table = uniform_matrix(T, seed + 1);
    std::vector<uint8_t> idx = index_matrix(n, k, density, T, seed + 2);
    Matrix S(idx.data(), n, k, k, table.data(), T, sblas::SblasTrans);      // S = idx^T : k x n
    if (S.NumRows() != k || S.NumCols() != n) { printf("CopyForm shape mismatch\n"); return 1; }
    std::vector<float> dense((size_t)k * n);
    S.CopyTo(dense.data(), n, sblas::SblasNoTrans);
    const int lda = k + pad_a, ldc = n + pad_c;
    std::vector<float> a = uniform_matrix((size_t)m * lda, seed + 3), c0 = uniform_matrix((size_t)m * ldc, seed + 4);
    std::vector<float> c = c0;
    S.AddMatMat(a.data(), m, lda, c.data(), ldc, alpha, beta);              // device path
    const ErrorMeasures e = compare_with_fp64(a, m, lda, c0, c, ldc, dense, k, n, alpha, beta);
    int bad = !(e.backward <= 1e-5) || !(e.normwise <= 1e-4);
    for (int i = 0; i < m && !bad; ++i)                                    // ldc padding untouched
        for (int j = n; j < ldc; ++j)
            if (memcmp(&c[(size_t)i * ldc + j], &c0[(size_t)i * ldc + j], 4) != 0) { bad = 1; break; }
    printf("%-4s m=%-4d n=%-5d k=%-5d density=%-6.4f alpha=%-4g beta=%-4g backward=%.2e normwise=%.2e\n",
           bad ? "FAIL" : "ok", m, n, k, density, alpha, beta, e.backward, e.normwise);
    return bad;
}

// Copy assignment after a product: `a` already has a device copy of its old contents registered under its
// address; `a = b` must drop it, so that the next AddMatMat multiplies by b's matrix (the reference class is
// correct under assignment, its implicit operator= copies the vectors).  A copy-constructed object too.
static int run_assignment_case() {
    const int T = 255, n = 300, k = 270, m = 5;
    std::vector<float> table = uniform_matrix(T, 901);
    std::vector<uint8_t> ia = index_matrix(n, k, 0.10, T, 902), ib = index_matrix(n, k, 0.15, T, 903);
    Matrix A(ia.data(), n, k, k, table.data(), T, sblas::SblasTrans), B(ib.data(), n, k, k, table.data(), T, sblas::SblasTrans);
    std::vector<float> x = uniform_matrix((size_t)m * k, 904), c0 = uniform_matrix((size_t)m * n, 905);
    std::vector<float> ca = c0, cb = c0, cc = c0, cd = c0;
    A.AddMatMat(x.data(), m, k, ca.data(), n, 1.3f, 2.0f);                 // registers A's device copy
    B.AddMatMat(x.data(), m, k, cb.data(), n, 1.3f, 2.0f);
    A = B;
    A.AddMatMat(x.data(), m, k, cc.data(), n, 1.3f, 2.0f);                 // must be B's product
    Matrix D(B);
    D.AddMatMat(x.data(), m, k, cd.data(), n, 1.3f, 2.0f);
    const int bad = memcmp(cc.data(), cb.data(), cb.size() * 4) != 0 || memcmp(cd.data(), cb.data(), cb.size() * 4) != 0 ||
                    memcmp(ca.data(), cb.data(), cb.size() * 4) == 0 || !(A == B);
    printf("%-4s assignment / copy construction after a product use the new contents\n", bad ? "FAIL" : "ok");
    return bad;
}

int main(int argc, char **argv) {
    int failures = 0;
    if (argc >= 5) {
        failures += run_case(atoi(argv[1]), atoi(argv[2]), atoi(argv[3]), atof(argv[4]), 1.0f, 1.0f, 0, 0, 1);
    } else {
        const int ms[] = {1, 2, 3, 7, 8, 9, 31, 117, 128};
        const float scal[][2] = {{1.0f, 1.0f}, {1.0f, 0.0f}, {1.3f, 2.0f}, {1.8f, 1.2f}, {0.0f, 0.5f}};
        int q = 0;
        for (int mi = 0; mi < 9; ++mi, ++q) {
            failures += run_case(ms[mi], 1023, 2047, 0.25, scal[q % 5][0], scal[q % 5][1], q % 3, q % 2, 100 + q);  // harness default shape
            failures += run_case(ms[mi], 700, 333, 0.02, scal[(q + 2) % 5][0], scal[(q + 2) % 5][1], 0, 5, 200 + q);
        }
        failures += run_case(1, 4096, 4096, 0.01, 1.0f, 1.0f, 0, 0, 11);      // BASELINE.json configs[0]
        failures += run_case(128, 4096, 4096, 0.10, 1.0f, 1.0f, 0, 0, 12);    // BASELINE.json configs[1]
        failures += run_assignment_case();
    }
    printf(failures ? "Parity Test fail... (%d case(s))\n" : "Parity Test success...\n", failures);
    return failures ? 1 : 0;
}
