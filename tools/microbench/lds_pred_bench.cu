// lds_pred_bench.cu -- does a predicated-off quarter warp of an LDS.128 still cost a shared-memory pipe cycle?
// 4 slots x 8 lanes; only slots < A are active.  Reports LDS.128 instructions per clk per SM and active wavefronts/clk.
#include <cstdio>
#include <cuda_runtime.h>

template <int U>
__global__ void __launch_bounds__(512) k(float *out, int iters, int rows, int active, long long *cycles) {
    extern __shared__ float4 smem[];
    const int lane = threadIdx.x & 31, slot = lane >> 3, c8 = lane & 7;
    for (int i = threadIdx.x; i < rows * 8; i += blockDim.x) smem[i] = make_float4(i, 1.f, 2.f, 3.f);
    __syncthreads();
    unsigned r = (threadIdx.x >> 5) * 7919u + slot * 104729u + 12345u;
    const float4 *addr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        r = r * 1664525u + 1013904223u;
        addr[u] = smem + (r >> 24) * 8 + c8;
    }
    float4 a4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool on = slot < active;
    long long t0 = clock64();
    float4 x[U];
#pragma unroll
    for (int u = 0; u < U; ++u) x[u] = a4;
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            asm volatile("{ .reg .pred p; setp.ne.u32 p, %5, 0; @p ld.shared.v4.f32 {%0,%1,%2,%3}, [%4]; }"
                         : "+f"(x[u].x), "+f"(x[u].y), "+f"(x[u].z), "+f"(x[u].w)
                         : "r"((unsigned)__cvta_generic_to_shared(addr[u])), "r"((unsigned)on));
        }
#pragma unroll
        for (int u = 0; u < U; ++u) { if (u & 1) a4.x += x[u].x; else a4.y += x[u].w; }
    }
    float acc = a4.x + a4.y + a4.z + a4.w;
    long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

int main() {
    const int sms = 148, iters = 2000, rows = 256, warps = 16;
    float *out; long long *cyc;
    cudaMalloc(&out, sms * 1024 * 4); cudaMalloc(&cyc, sms * 8);
    cudaFuncSetAttribute(k<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, rows * 128);
    for (int active = 1; active <= 4; ++active) {
        for (int rep = 0; rep < 2; ++rep) k<8><<<sms, warps * 32, rows * 128>>>(out, iters, rows, active, cyc);
        cudaDeviceSynchronize();
        long long h[148]; cudaMemcpy(h, cyc, sms * 8, cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < sms; ++i) avg += h[i]; avg /= sms;
        const double instrs = (double)warps * iters * 8;
        printf("active slots %d: %.3f LDS.128/clk/SM -> %.2f clk per LDS.128, active wavefronts/clk %.3f\n", active,
               instrs / avg, avg / instrs, instrs * active / avg);
    }
    return 0;
}
