// Pinned host buffers on this box: DMA copy (cudaMemcpyAsync) versus a kernel that reads the mapped
// pointer itself, per allocation.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o zero_copy_bench zero_copy_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <vector>

__global__ void copy_kernel(const float4 *__restrict__ src, float4 *__restrict__ dst, size_t n4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

static float time_it(cudaStream_t st, int reps, void (*fn)(void *), void *ctx) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) fn(ctx);
    cudaStreamSynchronize(st);
    cudaEventRecord(a, st);
    for (int i = 0; i < reps; ++i) fn(ctx);
    cudaEventRecord(b, st);
    cudaStreamSynchronize(st);
    float ms; cudaEventElapsedTime(&ms, a, b);
    return ms / reps * 1e3f;
}

struct Ctx { void *h, *hd, *d; size_t bytes; cudaStream_t st; int grid; };
static void do_dma_in(void *p) { Ctx *c = (Ctx *)p; cudaMemcpyAsync(c->d, c->h, c->bytes, cudaMemcpyHostToDevice, c->st); }
static void do_dma_out(void *p) { Ctx *c = (Ctx *)p; cudaMemcpyAsync(c->h, c->d, c->bytes, cudaMemcpyDeviceToHost, c->st); }
static void do_zc_in(void *p) { Ctx *c = (Ctx *)p; copy_kernel<<<c->grid, 256, 0, c->st>>>((const float4 *)c->hd, (float4 *)c->d, c->bytes / 16); }
static void do_zc_out(void *p) { Ctx *c = (Ctx *)p; copy_kernel<<<c->grid, 256, 0, c->st>>>((const float4 *)c->d, (float4 *)c->hd, c->bytes / 16); }

int main() {
    cudaStream_t st; cudaStreamCreate(&st);
    const size_t sizes[] = {16 << 10, 2 << 20, 2 << 20, 2 << 20, 32 << 20, 2 << 20, 32 << 20, 2 << 20};
    void *d; cudaMalloc(&d, 32 << 20);
    printf("%8s %10s %10s %12s %12s   (us; zero-copy kernel: grid 148*4 x 256, float4)\n", "KiB", "dma h2d", "dma d2h", "kernel h2d", "kernel d2h");
    for (size_t s : sizes) {
        Ctx c; c.bytes = s; c.st = st; c.d = d; c.grid = 148 * 4;
        cudaHostAlloc(&c.h, s, cudaHostAllocMapped);
        memset(c.h, 1, s);
        cudaHostGetDevicePointer(&c.hd, c.h, 0);
        const float a = time_it(st, 20, do_dma_in, &c), b = time_it(st, 20, do_dma_out, &c);
        const float z = time_it(st, 20, do_zc_in, &c), w = time_it(st, 20, do_zc_out, &c);
        printf("%8zu %10.1f %10.1f %12.1f %12.1f   h2d %.1f / %.1f GB/s\n", s >> 10, a, b, z, w, s / a / 1e3, s / z / 1e3);
    }
    return 0;
}
