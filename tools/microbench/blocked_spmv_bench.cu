// blocked_spmv_bench.cu -- is the column-blocked two-pass SpMV planned in DESIGN.md section 11 worth building?
// Both passes on synthetic data of the power-law case's size (2^27 entries, x of 2^23 floats, 296 row blocks x 256
// column blocks), without the format builder:
//   pass 1: one CTA per (column block, chunk of 65536 entries): x tile (32768 floats) into shared memory, then
//           p[e] = x_s[lcol[e]] * table[code[e]]  (2 + 1 bytes read, 4 written per entry, random shared-memory gathers)
//   pass 2: one CTA per row block: y tile in shared memory, the row block's 256 segments (sorted local rows, 2 + 4 bytes
//           per entry) added one after the other with a CTA barrier between segments; the next segment's entries are
//           loaded into registers before the current one is added.
// Prints microseconds per pass and GB/s.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o blocked_spmv_bench ...
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int CB = 32768, CHUNK = 65536, NCB = 256;

__device__ __forceinline__ uint32_t mix(uint32_t v) {
    v ^= v >> 16; v *= 0x7feb352du; v ^= v >> 15; v *= 0x846ca68bu; v ^= v >> 16;
    return v;
}

__global__ void fill_pass1(uint2 *lcol4, uint32_t *code4, size_t n4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t a = mix((uint32_t)i * 4u + 1u), b = mix((uint32_t)i * 4u + 2u), c = mix((uint32_t)i * 4u + 3u);
        lcol4[i] = make_uint2((a & 0x7fffu) | ((a >> 1) & 0x7fff0000u), (b & 0x7fffu) | ((b >> 1) & 0x7fff0000u));
        code4[i] = c;
    }
}

__global__ void fill_f32(float *p, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        p[i] = (float)(mix((uint32_t)i) & 0xffff) * (1.0f / 65536.0f) - 0.5f;
}

// sorted local rows inside every segment: entry t of a segment of length L gets row floor(t * RB / L)
__global__ void fill_lrow(uint16_t *lrow, const int *seg_ptr, int nseg, int RB) {
    const int s = blockIdx.x;
    if (s >= nseg) return;
    const int b = seg_ptr[s], e = seg_ptr[s + 1], L = e - b;
    for (int t = threadIdx.x; t < L; t += blockDim.x) lrow[b + t] = (uint16_t)(((long long)t * RB) / L);
}

template <int U>
__global__ void __launch_bounds__(1024, 1)
pass1_kernel(const uint2 *__restrict__ lcol4, const uint32_t *__restrict__ code4, const float *__restrict__ x,
             const float *__restrict__ table, float alpha, float4 *__restrict__ p4, int chunks_per_cb) {
    extern __shared__ float smem[];
    float *x_s = smem, *tab_s = smem + CB;
    const int cb = blockIdx.x / chunks_per_cb;
    const float4 *xg = reinterpret_cast<const float4 *>(x + (size_t)cb * CB);
    for (int i = threadIdx.x; i < CB / 4; i += 1024) reinterpret_cast<float4 *>(x_s)[i] = xg[i];
    if (threadIdx.x < 256) tab_s[threadIdx.x] = table[threadIdx.x] * alpha;
    __syncthreads();
    const size_t q0 = (size_t)blockIdx.x * (CHUNK / 4);
    for (int q = threadIdx.x; q < CHUNK / 4; q += 1024 * U) {      // U independent 12-byte loads in flight per thread
        uint2 lc[U];
        uint32_t cd[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { lc[u] = __ldcs(lcol4 + q0 + q + u * 1024); cd[u] = __ldcs(code4 + q0 + q + u * 1024); }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float4 p;
            p.x = x_s[lc[u].x & 0xffffu] * tab_s[cd[u] & 255u];
            p.y = x_s[lc[u].x >> 16] * tab_s[(cd[u] >> 8) & 255u];
            p.z = x_s[lc[u].y & 0xffffu] * tab_s[(cd[u] >> 16) & 255u];
            p.w = x_s[lc[u].y >> 16] * tab_s[cd[u] >> 24];
            __stcs(p4 + q0 + q + u * 1024, p);
        }
    }
}

// K = entries per thread and segment held in registers (segments of up to K * 1024 entries); D = segments loaded ahead
template <int K, int D>
__global__ void __launch_bounds__(1024, 1)
pass2_kernel(const uint16_t *__restrict__ lrow, const float *__restrict__ p, const int *__restrict__ seg_ptr,
             float *__restrict__ y, int RB, int n) {
    extern __shared__ float smem[];
    float *y_s = smem;
    int *sp = reinterpret_cast<int *>(smem + RB);
    const int rb = blockIdx.x;
    for (int i = threadIdx.x; i < RB; i += 1024) y_s[i] = 0.0f;
    for (int i = threadIdx.x; i <= NCB; i += 1024) sp[i] = seg_ptr[rb * NCB + i];
    __syncthreads();
    uint32_t r[D][K], rprev[D][K];
    float v[D][K];
    auto load = [&](int s, uint32_t (&rr)[K], uint32_t (&rp)[K], float (&vv)[K]) {
        const int b = sp[s], e = sp[s + 1];
#pragma unroll
        for (int q = 0; q < K; ++q) {
            const int i = b + threadIdx.x + q * 1024;
            rr[q] = 0xffffffffu; rp[q] = 0xfffffffeu; vv[q] = 0.0f;
            if (i < e) {
                rr[q] = lrow[i];
                vv[q] = __ldcs(p + i);
                if (i > b) rp[q] = lrow[i - 1];
            }
        }
    };
#pragma unroll
    for (int d = 0; d < D; ++d) load(d, r[d], rprev[d], v[d]);
    for (int s0 = 0; s0 < NCB; s0 += D) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const int s = s0 + d;
            uint32_t cr[K], cp[K];
            float cv[K];
#pragma unroll
            for (int q = 0; q < K; ++q) { cr[q] = r[d][q]; cp[q] = rprev[d][q]; cv[q] = v[d][q]; }
            if (s + D < NCB) load(s + D, r[d], rprev[d], v[d]);
#pragma unroll
            for (int q = 0; q < K; ++q) {
                if (cr[q] != 0xffffffffu) {
                    if (cr[q] != cp[q]) y_s[cr[q]] += cv[q];            // head of a run (runs have length 1 here)
                    else atomicAdd(&y_s[cr[q]], cv[q]);                  // stand-in for the in-run continuation
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < RB; i += 1024) {
        const size_t row = (size_t)rb * RB + i;
        if (row < (size_t)n) y[row] += y_s[i];
    }
}

int main() {
    const size_t nnz = (size_t)1 << 27;
    const int n = 1 << 23, k = 1 << 23;
    const int NRB = 296, RB = (n + NRB - 1) / NRB;
    const int nseg = NRB * NCB;
    uint2 *lcol4; uint32_t *code4; float *x, *table, *p, *y; uint16_t *lrow; int *seg_ptr;
    CK(cudaMalloc(&lcol4, nnz * 2)); CK(cudaMalloc(&code4, nnz)); CK(cudaMalloc(&x, (size_t)k * 4));
    CK(cudaMalloc(&table, 1024)); CK(cudaMalloc(&p, nnz * 4)); CK(cudaMalloc(&y, (size_t)n * 4));
    CK(cudaMalloc(&lrow, nnz * 2)); CK(cudaMalloc(&seg_ptr, (size_t)(nseg + 1) * 4));
    fill_pass1<<<148 * 8, 256>>>(lcol4, code4, nnz / 4);
    fill_f32<<<148 * 8, 256>>>(x, k);
    fill_f32<<<1, 256>>>(table, 256);
    CK(cudaMemset(y, 0, (size_t)n * 4));
    // segment lengths: nnz / nseg on average, +-12 % pseudo-random, last one takes the rest
    std::vector<int> sp(nseg + 1);
    {
        uint32_t st = 12345u;
        size_t pos = 0;
        const size_t avg = nnz / nseg;
        for (int s = 0; s < nseg; ++s) {
            sp[s] = (int)pos;
            st = st * 1664525u + 1013904223u;
            size_t len = avg - avg / 8 + (st >> 8) % (avg / 4);
            if (s == nseg - 1 || pos + len > nnz) len = nnz - pos;
            pos += len;
        }
        sp[nseg] = (int)pos;
        // the last segment took the remainder: move boundaries back until every segment has <= 2048 entries (K = 2)
        for (int s = nseg - 1; s >= 0 && sp[s + 1] - sp[s] > 2048; --s) sp[s] = sp[s + 1] - 2048;
    }
    CK(cudaMemcpy(seg_ptr, sp.data(), (size_t)(nseg + 1) * 4, cudaMemcpyHostToDevice));
    fill_lrow<<<nseg, 256>>>(lrow, seg_ptr, nseg, RB);
    CK(cudaDeviceSynchronize());
    int maxlen = 0;
    for (int s = 0; s < nseg; ++s) maxlen = sp[s + 1] - sp[s] > maxlen ? sp[s + 1] - sp[s] : maxlen;
    printf("entries %zu, segments %d (max %d entries), RB %d rows, covered %d\n", nnz, nseg, maxlen, RB, sp[nseg] - sp[0]);

    const int smem1 = (CB + 256) * 4, smem2 = RB * 4 + (NCB + 1) * 4;
    CK(cudaFuncSetAttribute(pass1_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem1));
    CK(cudaFuncSetAttribute(pass1_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem1));
    CK(cudaFuncSetAttribute(pass1_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem1));
    CK(cudaFuncSetAttribute((pass2_kernel<2, 1>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
    CK(cudaFuncSetAttribute((pass2_kernel<2, 2>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
    CK(cudaFuncSetAttribute((pass2_kernel<2, 4>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int items = (int)(nnz / CHUNK);
    auto time = [&](const char *what, auto launch, double bytes) {
        for (int w = 0; w < 2; ++w) launch();
        CK(cudaDeviceSynchronize());
        float best = 1e9f, sum = 0.0f;
        for (int it = 0; it < 8; ++it) {
            CK(cudaEventRecord(e0));
            launch();
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            best = ms < best ? ms : best; sum += ms;
        }
        CK(cudaGetLastError());
        printf("%-28s mean %8.1f us  min %8.1f us  %7.1f GB/s\n", what, sum / 8 * 1e3, best * 1e3, bytes / (sum / 8 * 1e-3) * 1e-9);
    };
    const int cpc = items / (k / CB);
    const double b1 = (double)nnz * 7 + (double)items * CB * 4;
    time("pass 1, 2 loads in flight", [&] { pass1_kernel<2><<<items, 1024, smem1>>>(lcol4, code4, x, table, 1.0f, (float4 *)p, cpc); }, b1);
    time("pass 1, 4 loads in flight", [&] { pass1_kernel<4><<<items, 1024, smem1>>>(lcol4, code4, x, table, 1.0f, (float4 *)p, cpc); }, b1);
    time("pass 1, 8 loads in flight", [&] { pass1_kernel<8><<<items, 1024, smem1>>>(lcol4, code4, x, table, 1.0f, (float4 *)p, cpc); }, b1);
    time("pass 2, 1 segment ahead", [&] { pass2_kernel<2, 1><<<NRB, 1024, smem2>>>(lrow, p, seg_ptr, y, RB, n); }, (double)nnz * 6);
    time("pass 2, 2 segments ahead", [&] { pass2_kernel<2, 2><<<NRB, 1024, smem2>>>(lrow, p, seg_ptr, y, RB, n); }, (double)nnz * 6);
    time("pass 2, 4 segments ahead", [&] { pass2_kernel<2, 4><<<NRB, 1024, smem2>>>(lrow, p, seg_ptr, y, RB, n); }, (double)nnz * 6);
    return 0;
}
