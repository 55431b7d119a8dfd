// lds_bench.cu -- what does the shared-memory pipe of one SM deliver for the SpMM access pattern?
// Each warp: 4 slots x 8 lanes, every slot reads one 128-byte row (LDS.128 per lane) per step, U steps
// in flight, followed by 4 FFMA per loaded float4.  One CTA per SM, W warps.  Prints bytes/clk/SM.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

template <int U, bool FMA>
__global__ void lds_kernel(float *out, int iters, int rows, long long *cycles) {
    extern __shared__ float4 smem[];
    const int lane = threadIdx.x & 31, slot = lane >> 3, c8 = lane & 7;
    for (int i = threadIdx.x; i < rows * 8; i += blockDim.x) smem[i] = make_float4(i, 1.f, 2.f, 3.f);
    __syncthreads();
    unsigned r = (threadIdx.x >> 5) * 7919u + slot * 104729u + 12345u;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        float4 x[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            r = r * 1664525u + 1013904223u;
            const int row = (r >> 8) % rows;
            x[u] = smem[row * 8 + c8];
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (FMA) {
                acc.x = fmaf(x[u].x, 1.0001f, acc.x); acc.y = fmaf(x[u].y, 1.0001f, acc.y);
                acc.z = fmaf(x[u].z, 1.0001f, acc.z); acc.w = fmaf(x[u].w, 1.0001f, acc.w);
            } else {
                acc.x += x[u].x;
            }
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc.x + acc.y + acc.z + acc.w;
}

template <int U, bool FMA>
void run(int warps, int iters, int rows) {
    int sms = 148;
    float *out; long long *cyc;
    cudaMalloc(&out, sms * 1024 * 4); cudaMalloc(&cyc, sms * 8);
    const size_t smem = (size_t)rows * 128;
    cudaFuncSetAttribute(lds_kernel<U, FMA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    lds_kernel<U, FMA><<<sms, warps * 32, smem>>>(out, iters, rows, cyc);
    cudaDeviceSynchronize();
    lds_kernel<U, FMA><<<sms, warps * 32, smem>>>(out, iters, rows, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sms * 8, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < sms; ++i) avg += h[i]; avg /= sms;
    const double bytes = (double)warps * iters * U * 512.0;
    printf("U=%2d fma=%d warps=%2d rows=%4d : %.1f B/clk/SM  (%.3f wavefronts/clk)  %s\n", U, (int)FMA, warps, rows,
           bytes / avg, bytes / avg / 128.0, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    const int iters = 2000;
    for (int warps : {4, 8, 14, 16, 28, 32}) {
        run<4, true>(warps, iters, 255);
        run<8, true>(warps, iters, 255);
        run<16, true>(warps, iters, 255);
        run<8, false>(warps, iters, 255);
    }
    run<8, true>(16, iters, 1);      // all slots on one row: broadcast
    run<8, true>(16, iters, 765);    // 3 stages worth of rows
    return 0;
}
