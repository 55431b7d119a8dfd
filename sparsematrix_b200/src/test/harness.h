// harness.h -- shared pieces of the C++ parity / benchmark harness (src/test).
//
// Mirrors what the reference's harness does (reference src/test/blas_test.h): a dense uint8 index
// matrix n x k with a fraction of "zero" indices, a 255-entry codebook, CopyForm(..., SblasTrans),
// C = A * S against a dense product on the CopyTo-expanded matrix -- with three differences the
// survey calls for: seeded generators (std::mt19937_64 instead of rand()/
// srand(time)), an fp64 dense product as the yardstick with the backward-error measure, and a
// non-zero exit code on failure.
#ifndef SBLAS_B200_TEST_HARNESS_H_
#define SBLAS_B200_TEST_HARNESS_H_

#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <random>
#include <string>
#include <vector>

#include "sparse-matrix.h"

namespace harness {

typedef sblas::SparseMatrix<uint8, uint8, float> Matrix;

inline std::vector<float> uniform_matrix(size_t count, uint64_t seed) {
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<float> dist(-1.0f, 1.0f);
    std::vector<float> v(count);
    for (size_t i = 0; i < count; ++i) v[i] = dist(rng);
    return v;
}

// n x k index matrix, each element stored with probability `density` (index uniform in [0, T)), else 255
inline std::vector<uint8_t> index_matrix(int n, int k, double density, int T, uint64_t seed) {
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> u(0.0, 1.0);
    std::uniform_int_distribution<int> pick(0, T - 1);
    std::vector<uint8_t> idx((size_t)n * k);
    for (size_t i = 0; i < idx.size(); ++i) idx[i] = u(rng) < density ? (uint8_t)pick(rng) : (uint8_t)255;
    return idx;
}

struct ErrorMeasures {
    double backward;   // max |y - y64| / (|beta c| + |alpha| sum |a||s|)   -- the <= 1e-5 contract
    double normwise;   // max |y - y64| / max |y64|
};

// fp64 dense product on the CopyTo-expanded S (dense_s: k x n row-major): y64 = beta*c0 + alpha*A*S
inline ErrorMeasures compare_with_fp64(const std::vector<float> &a, int m, int lda, const std::vector<float> &c0,
                                       const std::vector<float> &c, int ldc, const std::vector<float> &dense_s, int k,
                                       int n, float alpha, float beta) {
    ErrorMeasures e = {0.0, 0.0};
    double ymax = 0.0, dmax = 0.0;
    std::vector<double> y((size_t)n), bound((size_t)n);
    for (int i = 0; i < m; ++i) {
        for (int j = 0; j < n; ++j) {
            y[j] = (double)beta * c0[(size_t)i * ldc + j];
            bound[j] = fabs(y[j]);
        }
        for (int r = 0; r < k; ++r) {
            const double av = (double)alpha * a[(size_t)i * lda + r];
            if (av == 0.0) continue;
            const float *srow = &dense_s[(size_t)r * n];
            for (int j = 0; j < n; ++j) {
                const double p = av * srow[j];
                y[j] += p;
                bound[j] += fabs(p);
            }
        }
        for (int j = 0; j < n; ++j) {
            const double d = fabs((double)c[(size_t)i * ldc + j] - y[j]);
            if (bound[j] > 0 && d / bound[j] > e.backward) e.backward = d / bound[j];
            if (bound[j] == 0 && d != 0) e.backward = INFINITY;
            if (d > dmax) dmax = d;
            if (fabs(y[j]) > ymax) ymax = fabs(y[j]);
        }
    }
    e.normwise = ymax > 0 ? dmax / ymax : dmax;
    return e;
}

// "a" or "a:b" -> doubling sweep a, 2a, ... <= b (the reference's StepList, blas_test.h)
inline std::vector<int> parse_steps(const char *arg) {
    std::vector<int> v;
    int a = atoi(arg), b = a;
    const char *colon = strchr(arg, ':');
    if (colon) b = atoi(colon + 1);
    if (a < 1) a = 1;
    for (int x = a; x <= b; x *= 2) v.push_back(x);
    if (v.empty()) v.push_back(a);
    return v;
}

}  // namespace harness

#endif
