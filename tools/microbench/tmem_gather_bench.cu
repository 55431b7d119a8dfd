// tmem_gather_bench.cu -- can tensor memory serve as a second operand store for the SpMM gather?
//
// The SpMM tile kernel is bound by the shared-memory pipe: every non-zero needs its own 4*m bytes of
// X and one SM delivers 128 B per clock from shared memory (lds_bench.cu).  Tensor memory (256 KB per
// SM, 128 lanes x 512 columns x 32 bit) is read with tcgen05.ld by ordinary warps -- normally for a
// GEMM epilogue.  With lane = dense column and column = X row (x4: four dense columns per thread and
// row), a warp-uniform, data-dependent column address is exactly the gather the SpMM needs.  This
// benchmark measures how many bytes per clock and SM such loads deliver, alone and next to LDS.128
// gathers from shared memory.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_gather_bench tmem_gather_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

__device__ __forceinline__ void tmem_ld_x4(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_x1(uint32_t taddr, uint32_t &r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(taddr));
}
__device__ __forceinline__ uint32_t tmem_ld_x8(uint32_t taddr) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
    return r[0] ^ r[1] ^ r[2] ^ r[3] ^ r[4] ^ r[5] ^ r[6] ^ r[7];
}
__device__ __forceinline__ uint32_t tmem_ld_x16(uint32_t taddr) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                   "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) : "r"(taddr));
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) x ^= r[i];
    return x;
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// MODE 0: tcgen05.ld x4 only, 1: LDS.128 only, 2: both in every step, 3: tcgen05.ld x1 only, 4: x8, 5: x16 (the xor of the
// registers is formed after the wait; x8 / x16 read 8 / 16 consecutive columns, i.e. need m >= 1024 / 2048 to be one X row)
template <int MODE>
__global__ void __launch_bounds__(512, 1)
gather_kernel(int iters, unsigned long long *cycles, uint32_t *sink) {
    extern __shared__ __align__(16) uint8_t smem[];       // 128 rows x 512 B of "X" for the LDS leg
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_base_s)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    for (int i = threadIdx.x; i < 128 * 128; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = i;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tbase = tmem_base_s + ((uint32_t)((warp & 3) * 32) << 16);
    if (warp < 4) {                                       // define the contents: column c of lane l = c * 128 + l
        for (int c = 0; c < 512; c += 4) {
            const uint32_t v = c * 128 + (warp & 3) * 32 + lane;
            asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(tbase + c), "r"(v), "r"(v + 128), "r"(v + 256), "r"(v + 384));
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    uint32_t acc = 0, rng = 12345u + warp * 977u;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem) + (lane & 7) * 16;   // 8 lanes x 16 B = one 128-byte row segment per slot
    const int slot = lane >> 3;
    const unsigned long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        uint32_t r[8][4];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            rng = rng * 1664525u + 1013904223u;           // warp-uniform "X row" of this step
            const uint32_t row = (rng >> 8) & 127u;
            if (MODE == 0 || MODE == 2) tmem_ld_x4(tbase + row * 4, r[u]);
            if (MODE == 3) { tmem_ld_x1(tbase + row * 4, r[u][0]); r[u][1] = r[u][2] = r[u][3] = 0; }
            if (MODE == 4) { r[u][0] = tmem_ld_x8(tbase + (row & 63u) * 8); r[u][1] = r[u][2] = r[u][3] = 0; }
            if (MODE == 5) { r[u][0] = tmem_ld_x16(tbase + (row & 31u) * 16); r[u][1] = r[u][2] = r[u][3] = 0; }
            if (MODE == 1 || MODE == 2) {
                const uint32_t srow = (row + slot * 31u) & 127u;   // four slots, four different rows, like the tile kernel
                uint32_t a, b, c, d;
                asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(sbase + srow * 512));
                if (MODE == 1) { r[u][0] = a; r[u][1] = b; r[u][2] = c; r[u][3] = d; }
                else acc += a ^ b ^ c ^ d;
            }
        }
        if (MODE != 1) tmem_wait_ld();
#pragma unroll
        for (int u = 0; u < 8; ++u) acc += r[u][0] ^ r[u][1] ^ r[u][2] ^ r[u][3];
    }
    const unsigned long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base_s));
}

template <int MODE>
static int run(const char *name, int warps, int iters, double bytes_tmem, double bytes_lds) {
    unsigned long long *d_cycles; uint32_t *d_sink;
    CK(cudaMalloc(&d_cycles, 148 * 8));
    CK(cudaMalloc(&d_sink, 148 * 512 * 4));
    CK(cudaFuncSetAttribute(gather_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    gather_kernel<MODE><<<148, warps * 32, 65536>>>(iters, d_cycles, d_sink);
    gather_kernel<MODE><<<148, warps * 32, 65536>>>(iters, d_cycles, d_sink);
    CK(cudaDeviceSynchronize());
    unsigned long long h[148];
    CK(cudaMemcpy(h, d_cycles, sizeof(h), cudaMemcpyDeviceToHost));
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += (double)h[i] / 148;
    const double steps = (double)iters * 8 * warps;       // warp-level loads per SM
    printf("%-34s %2d warps: %8.0f clk, %6.2f clk per warp-load, tmem %7.1f B/clk/SM, smem %7.1f B/clk/SM\n", name, warps, avg,
           avg / steps, bytes_tmem * steps / avg, bytes_lds * steps / avg);
    cudaFree(d_cycles); cudaFree(d_sink);
    return 0;
}

int main() {
    const int iters = 4000;
    for (int warps : {4, 8, 16}) {
        if (run<1>("LDS.128, 4 rows per warp-load", warps, iters, 0, 512)) return 1;
        if (run<0>("tcgen05.ld 32x32b.x4", warps, iters, 512, 0)) return 1;
        if (run<3>("tcgen05.ld 32x32b.x1", warps, iters, 128, 0)) return 1;
        if (run<2>("both per step", warps, iters, 512, 512)) return 1;
        if (run<4>("tcgen05.ld 32x32b.x8", warps, iters, 1024, 0)) return 1;
        if (run<5>("tcgen05.ld 32x32b.x16", warps, iters, 2048, 0)) return 1;
    }
    return 0;
}
