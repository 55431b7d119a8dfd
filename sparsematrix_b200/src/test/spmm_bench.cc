// spmm_bench.cc -- timed harness in the style of the reference's blas_test
// (src/test/blas_test.cc:31-47: "blas_test [m[:mEnd]] [n[:nEnd]] [k[:kEnd]] ..." doubling sweeps,
// markdown table of milliseconds), extended with what SURVEY.md 8(d) asks of a GPU build:
// device-resident operands, warm-up, CUDA-event timing over many launches, cold (L2 flushed) and
// warm figures, GFLOP/s and effective GB/s, plus the host-pointer call (copies included).
//
//   spmm_bench [m[:mEnd]] [n[:nEnd]] [k[:kEnd]] [density] [iters]
// defaults: 128 4096 4096 0.10 50   (BASELINE.json configs[1])
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>

#include "harness.h"
#include "sblas_b200.h"

using namespace harness;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } \
    } while (0)

struct Row { int m, n, k; long long nnz; double cold_us, warm_us, host_us, pinned_us, gflops, gbs; };

static Row bench_one(int m, int n, int k, double density, int iters) {
    const int T = 255;
    std::vector<float> table = uniform_matrix(T, 1);
    std::vector<uint8_t> idx = index_matrix(n, k, density, T, 12);
    Matrix S(idx.data(), n, k, k, table.data(), T, sblas::SblasTrans);
    sblas_b200_matrix *dev = S.DeviceMatrix();
    if (!dev) exit(2);
    sblas_b200_info info;
    sblas_b200_get_info(dev, &info);
    std::vector<float> a = uniform_matrix((size_t)m * k, 2), c((size_t)m * n, 0.0f);
    float *d_a, *d_c;
    char *d_flush;
    const size_t flush_bytes = 256u << 20;
    CK(cudaMalloc(&d_a, a.size() * 4));
    CK(cudaMalloc(&d_c, c.size() * 4));
    CK(cudaMalloc(&d_flush, flush_bytes));
    CK(cudaMemcpy(d_a, a.data(), a.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(d_c, 0, c.size() * 4));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 5; ++i) sblas_b200_addmatmat_dev(dev, d_a, m, k, d_c, n, 1.0f, 1.0f, st);   // alpha = beta = 1 (blas_test.h:313)
    CK(cudaStreamSynchronize(st));
    double cold_ms = 0.0;
    for (int i = 0; i < iters; ++i) {
        CK(cudaMemsetAsync(d_flush, i, flush_bytes, st));          // evict L2 (126 MB) between launches
        CK(cudaEventRecord(e0, st));
        if (sblas_b200_addmatmat_dev(dev, d_a, m, k, d_c, n, 1.0f, 1.0f, st) != SBLAS_B200_OK) { printf("%s\n", sblas_b200_last_error()); exit(2); }
        CK(cudaEventRecord(e1, st));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        cold_ms += ms;
    }
    CK(cudaEventRecord(e0, st));
    for (int i = 0; i < iters; ++i) sblas_b200_addmatmat_dev(dev, d_a, m, k, d_c, n, 1.0f, 1.0f, st);
    CK(cudaEventRecord(e1, st));
    CK(cudaEventSynchronize(e1));
    float warm_ms;
    CK(cudaEventElapsedTime(&warm_ms, e0, e1));
    // the reference's own call shape: host pointers in, host pointers out
    const int host_iters = std::max(3, iters / 5);
    S.AddMatMat(a.data(), m, k, c.data(), n, 1.0f, 1.0f);
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < host_iters; ++i) S.AddMatMat(a.data(), m, k, c.data(), n, 1.0f, 1.0f);
    const double host_us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / host_iters;
    // the same call with a and c from the library's allocator (page-locked + mapped): copies overlap the product
    float *pa = nullptr, *pc = nullptr;
    if (sblas_b200_host_alloc((int64_t)a.size() * 4, (void **)&pa) != SBLAS_B200_OK ||
        sblas_b200_host_alloc((int64_t)c.size() * 4, (void **)&pc) != SBLAS_B200_OK) { printf("%s\n", sblas_b200_last_error()); exit(2); }
    std::copy(a.begin(), a.end(), pa);
    std::fill(pc, pc + c.size(), 0.0f);
    S.AddMatMat(pa, m, k, pc, n, 1.0f, 1.0f);
    const auto t1 = std::chrono::steady_clock::now();
    for (int i = 0; i < host_iters; ++i) S.AddMatMat(pa, m, k, pc, n, 1.0f, 1.0f);
    const double pinned_us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t1).count() / host_iters;
    sblas_b200_host_free(pa); sblas_b200_host_free(pc);
    Row r;
    r.pinned_us = pinned_us;
    r.m = m; r.n = n; r.k = k; r.nnz = info.nnz;
    r.cold_us = cold_ms * 1e3 / iters;
    r.warm_us = warm_ms * 1e3 / iters;
    r.host_us = host_us;
    r.gflops = 2.0 * info.nnz * m / (r.cold_us * 1e-6) / 1e9;
    r.gbs = (5.0 * info.nnz + 4.0 * (n + 1) + 4.0 * k * m + 8.0 * (double)n * m) / (r.cold_us * 1e-6) / 1e9;
    cudaFree(d_a); cudaFree(d_c); cudaFree(d_flush);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
    return r;
}

int main(int argc, char **argv) {
    std::vector<int> ms = parse_steps(argc > 1 ? argv[1] : "128"), ns = parse_steps(argc > 2 ? argv[2] : "4096"),
                     ks = parse_steps(argc > 3 ? argv[3] : "4096");
    const double density = argc > 4 ? atof(argv[4]) : 0.10;
    const int iters = argc > 5 ? atoi(argv[5]) : 50;
    std::vector<Row> rows;
    for (int m : ms) for (int n : ns) for (int k : ks) rows.push_back(bench_one(m, n, k, density, iters));
    printf("| m | n | k | nnz | cold us | warm us | host-pointer us (malloc) | host-pointer us (sblas_b200_host_alloc) | GFLOP/s (cold) | GB/s (cold) |\n|---|---|---|---|---|---|---|---|---|---|\n");
    for (const Row &r : rows)
        printf("| %d | %d | %d | %lld | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f |\n", r.m, r.n, r.k, r.nnz, r.cold_us, r.warm_us,
               r.host_us, r.pinned_us, r.gflops, r.gbs);
    for (const Row &r : rows)
        printf("{\"m\": %d, \"n\": %d, \"k\": %d, \"density\": %g, \"nnz\": %lld, \"cold_us\": %.2f, \"warm_us\": %.2f, "
               "\"host_pointer_us\": %.2f, \"host_pointer_pinned_us\": %.2f, \"gflops\": %.2f, \"effective_gbs\": %.2f}\n",
               r.m, r.n, r.k, density, r.nnz, r.cold_us, r.warm_us, r.host_us, r.pinned_us, r.gflops, r.gbs);
    return 0;
}
