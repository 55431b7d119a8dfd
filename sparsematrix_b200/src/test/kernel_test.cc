// kernel_test.cc -- the reference's kernel unit test (src/sparse/kernel_test.cc:27-57 there: sblas_trans_kernel
// against a naive transpose, bit for bit, 1023 x 511 with lda 512 and ldsa 1024) through the device-backed free
// functions of kernel.h, extended to the other free functions: the beta scale (exact) and the four panel
// formulations (against an fp64 loop over the decoded panel, backward error <= 1e-5).  Exit code != 0 on failure.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <random>
#include <vector>

#include "kernel.h"

static std::vector<float> uniform(size_t count, uint64_t seed) {
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<float> dist(-1.0f, 1.0f);
    std::vector<float> v(count);
    for (float &x : v) x = dist(rng);
    return v;
}

int main() {
    int failures = 0;
    {   // transpose, the reference's own shape
        const int m = 1023, n = 511, lda = 512, ldsa = 1024;
        std::vector<float> a = uniform((size_t)1024 * 512, 1), want((size_t)512 * 1024, -7.0f), got = want;
        for (int i = 0; i < m; ++i)
            for (int j = 0; j < n; ++j) want[(size_t)j * ldsa + i] = a[(size_t)i * lda + j];
        sblas_trans_kernel(a.data(), m, n, lda, got.data(), ldsa);
        const bool ok = memcmp(want.data(), got.data(), want.size() * 4) == 0;   // including everything that must stay untouched
        printf("%-4s sblas_trans_kernel 1023 x 511\n", ok ? "ok" : "FAIL");
        failures += !ok;
    }
    {   // beta scale
        const int m = 37, n = 301, ldc = 320;
        std::vector<float> c = uniform((size_t)m * ldc, 2), want = c;
        for (int i = 0; i < m; ++i)
            for (int j = 0; j < n; ++j) want[(size_t)i * ldc + j] *= 1.7f;
        sblas_beta_operation_kernel(c.data(), m, n, ldc, 1.7f);
        const bool ok = memcmp(want.data(), c.data(), c.size() * 4) == 0;
        printf("%-4s sblas_beta_operation_kernel 37 x 301\n", ok ? "ok" : "FAIL");
        failures += !ok;
    }
    {   // one panel, four formulations
        const int m = 9, n = 200, k = 333, T = 255;
        std::vector<float> table = uniform(T + 1, 3);
        table[T] = 0.0f;
        // panel k x n at 15 % density as the reference's byte streams (sparse-matrix.cc:35-60), fillers included
        std::mt19937_64 rng(4);
        std::vector<uint8_t> pos, val;
        std::vector<double> dense((size_t)k * n, 0.0);
        int prev = 0;
        for (int r = 0; r < k; ++r)
            for (int j = 0; j < n; ++j) {
                if (rng() % 100 >= 15) continue;
                const int code = (int)(rng() % T), lin = r * 256 + j;
                int d = lin - prev;
                while (d > 255) { pos.push_back(255); val.push_back((uint8_t)T); d -= 255; }
                pos.push_back((uint8_t)d); val.push_back((uint8_t)code);
                prev = lin;
                dense[(size_t)r * n + j] = table[code];
            }
        std::vector<float> a = uniform((size_t)m * k, 5), c0 = uniform((size_t)m * n, 6);
        const float alpha = 1.3f;
        std::vector<double> y((size_t)m * n), bound((size_t)m * n);
        for (int i = 0; i < m; ++i)
            for (int j = 0; j < n; ++j) {
                double s = c0[(size_t)i * n + j], b = fabs(s);
                for (int r = 0; r < k; ++r) {
                    const double p = (double)alpha * a[(size_t)i * k + r] * dense[(size_t)r * n + j];
                    s += p; b += fabs(p);
                }
                y[(size_t)i * n + j] = s; bound[(size_t)i * n + j] = b;
            }
        for (int which = 0; which < 4; ++which) {
            std::vector<float> c = c0;
            if (which < 2) {
                if (which == 0) sblas_kernel_operation<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(m, n, k, a.data(), k, c.data(), n, alpha, pos.data(), val.data(), (int)pos.size(), table.data(), T);
                else sblas_kernel_operation_naive<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(m, n, k, a.data(), k, c.data(), n, alpha, pos.data(), val.data(), (int)pos.size(), table.data(), T);
            } else {
                const int ld = 16;                                   // aligned_m of sparse-matrix.cc:155
                std::vector<float> at((size_t)k * ld, 0.0f), ct((size_t)n * ld, 0.0f);
                for (int i = 0; i < m; ++i) {
                    for (int r = 0; r < k; ++r) at[(size_t)r * ld + i] = a[(size_t)i * k + r];
                    for (int j = 0; j < n; ++j) ct[(size_t)j * ld + i] = c0[(size_t)i * n + j];
                }
                if (which == 2) sblas_kernel_operation_trans<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(m, n, k, at.data(), ld, ct.data(), ld, alpha, pos.data(), val.data(), (int)pos.size(), table.data(), T);
                else sblas_kernel_operation_trans_ex<uint8_t, uint8_t, float, SBLAS_BLOCK_COL_SHIFT>(m, n, k, at.data(), ld, ct.data(), ld, alpha, pos.data(), val.data(), (int)pos.size(), table.data(), T);
                for (int i = 0; i < m; ++i)
                    for (int j = 0; j < n; ++j) c[(size_t)i * n + j] = ct[(size_t)j * ld + i];
            }
            double worst = 0.0;
            for (size_t i = 0; i < c.size(); ++i) worst = fmax(worst, fabs((double)c[i] - y[i]) / bound[i]);
            const bool ok = worst <= 1e-5;
            printf("%-4s panel formulation %d: backward error %.2e\n", ok ? "ok" : "FAIL", which, worst);
            failures += !ok;
        }
    }
    printf(failures ? "Kernel Test fail...\n" : "Kernel Test success...\n");
    return failures ? -1 : 0;
}
