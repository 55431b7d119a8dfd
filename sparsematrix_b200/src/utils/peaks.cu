// peaks.cu -- the compute peak the roofline divides by, measured instead of assumed ("P_fp32 = measured
// FMA micro-benchmark, else 2*128*SMs*clock").  Register-resident FMA chains, enough of them to keep every scheduler busy.
#include "device_matrix.h"

namespace sblas_b200 {
namespace {

constexpr int kChains = 8, kIters = 4096;

__global__ void __launch_bounds__(1024)
fma_peak_kernel(float *__restrict__ out, float a, float b) {
    float v[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) v[c] = (float)(threadIdx.x + c);
    for (int it = 0; it < kIters; ++it) {
#pragma unroll
        for (int c = 0; c < kChains; ++c) v[c] = fmaf(v[c], a, b);
    }
    float s = 0.0f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += v[c];
    if (s == 12345.678f) out[0] = s;                         // keeps the chains alive
}

}  // namespace
}  // namespace sblas_b200

using namespace sblas_b200;

extern "C" int sblas_b200_measure_fp32_peak(int32_t device, double *tflops) {
    if (!tflops) return fail(SBLAS_B200_EINVAL, "tflops is null");
    *tflops = 0.0;
    DeviceScope scope(device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    int sms = 0;
    SB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    float *d_out = nullptr;
    SB_CUDA(cudaMalloc((void **)&d_out, 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = sms * 2 * 8;                            // two CTAs of 1024 threads per SM, eight waves
    double best_ms = 1e30;
    cudaError_t err = cudaSuccess;
    for (int rep = 0; rep < 5 && err == cudaSuccess; ++rep) {
        cudaEventRecord(e0, 0);
        SB_LAUNCH(fma_peak_kernel, grid, 1024, 0, 0, d_out, 0.999f, 0.001f);
        cudaEventRecord(e1, 0);
        err = cudaEventSynchronize(e1);
        float ms = 0.0f;
        if (err == cudaSuccess) err = cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms > 0.0f && ms < best_ms) best_ms = ms;     // the first launch warms up
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    if (err != cudaSuccess) return cuda_fail(err, "fp32 peak measurement", __FILE__, __LINE__);
    *tflops = 2.0 * (double)grid * 1024.0 * kChains * kIters / (best_ms * 1e-3) / 1e12;
    return SBLAS_B200_OK;
}
