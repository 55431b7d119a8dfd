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

constexpr int CHUNK = 65536, NCB = 256;

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

// U = independent 12-byte loads in flight per thread; MODE 0 = the kernel, 1 = without the stores (ablation), 2 = without the
// shared-memory gathers (ablation); CBT = columns per block (x tile), NT = threads per CTA
template <int U, int MODE, int CBT, int NT>
__global__ void __launch_bounds__(NT, 1024 / NT)
pass1_kernel(const uint2 *__restrict__ lcol4, const uint32_t *__restrict__ code4, const float *__restrict__ x,
             const float *__restrict__ table, float alpha, float4 *__restrict__ p4, int chunks_per_cb) {
    extern __shared__ float smem[];
    float *x_s = smem, *tab_s = smem + CBT;
    const int cb = blockIdx.x / chunks_per_cb;
    const float4 *xg = reinterpret_cast<const float4 *>(x + (size_t)cb * CBT);
    for (int i = threadIdx.x; i < CBT / 4; i += NT) reinterpret_cast<float4 *>(x_s)[i] = xg[i];
    if (threadIdx.x < 256) tab_s[threadIdx.x] = table[threadIdx.x] * alpha;
    __syncthreads();
    const size_t q0 = (size_t)blockIdx.x * (CHUNK / 4);
    float sink = 0.0f;
    for (int q = threadIdx.x; q < CHUNK / 4; q += NT * U) {
        uint2 lc[U];
        uint32_t cd[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { lc[u] = __ldcs(lcol4 + q0 + q + u * NT); cd[u] = __ldcs(code4 + q0 + q + u * NT); }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float4 p;
            if (MODE == 2) {
                p = make_float4(__uint_as_float(lc[u].x), __uint_as_float(lc[u].y), __uint_as_float(cd[u]), 1.0f);
            } else {
                p.x = x_s[lc[u].x & (CBT - 1)] * tab_s[cd[u] & 255u];
                p.y = x_s[(lc[u].x >> 16) & (CBT - 1)] * tab_s[(cd[u] >> 8) & 255u];
                p.z = x_s[lc[u].y & (CBT - 1)] * tab_s[(cd[u] >> 16) & 255u];
                p.w = x_s[(lc[u].y >> 16) & (CBT - 1)] * tab_s[cd[u] >> 24];
            }
            if (MODE == 1) sink += p.x + p.y + p.z + p.w;
            else __stcs(p4 + q0 + q + u * NT, p);
        }
    }
    if (MODE == 1 && sink == 12345.678f) p4[0].x = sink;
}

// K = entries per thread and segment held in registers (segments of up to K * 1024 entries); D = segments loaded ahead
// MODE 0 = the kernel, 1 = without the barriers (ablation, wrong sums), 2 = without the shared-memory updates (ablation)
template <int K, int D, int MODE, int NT>
__global__ void __launch_bounds__(NT, 1024 / NT)
pass2_kernel(const uint16_t *__restrict__ lrow, const float *__restrict__ p, const int *__restrict__ seg_ptr,
             float *__restrict__ y, int RB, int n) {
    extern __shared__ float smem[];
    float *y_s = smem;
    int *sp = reinterpret_cast<int *>(smem + RB);
    const int rb = blockIdx.x;
    for (int i = threadIdx.x; i < RB; i += NT) y_s[i] = 0.0f;
    for (int i = threadIdx.x; i <= NCB; i += NT) sp[i] = seg_ptr[rb * NCB + i];
    __syncthreads();
    uint32_t r[D][K], rprev[D][K];
    float v[D][K];
    auto load = [&](int s, uint32_t (&rr)[K], uint32_t (&rp)[K], float (&vv)[K]) {
        const int b = sp[s], e = sp[s + 1];
#pragma unroll
        for (int q = 0; q < K; ++q) {
            const int i = b + threadIdx.x + q * NT;
            rr[q] = 0xffffffffu; rp[q] = 0xfffffffeu; vv[q] = 0.0f;
            if (i < e) {
                rr[q] = lrow[i];
                vv[q] = __ldcs(p + i);
                if (i > b) rp[q] = lrow[i - 1];
            }
        }
    };
    float sink = 0.0f;
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
                    if (MODE == 2) sink += cv[q] + __uint_as_float(cr[q] ^ cp[q]);
                    else if (cr[q] != cp[q]) y_s[cr[q]] += cv[q];       // head of a run (runs have length 1 here)
                    else atomicAdd(&y_s[cr[q]], cv[q]);                  // stand-in for the in-run continuation
                }
            }
            if (MODE != 1) __syncthreads();
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RB; i += NT) {
        const size_t row = (size_t)rb * RB + i;
        if (row < (size_t)n) y[row] += y_s[i] + (sink == 12345.678f ? 1.0f : 0.0f);
    }
}

// pass 1, persistent: G CTAs, CTA c takes the contiguous entry range [c, c + 1) * nnz / G and refills its x tile only where
// the range crosses into the next column block (every column block holds per_cb entries in the synthetic data).
template <int U, int CBT, int NT>
__global__ void __launch_bounds__(NT, 1024 / NT)
pass1p_kernel(const uint2 *__restrict__ lcol4, const uint32_t *__restrict__ code4, const float *__restrict__ x,
              const float *__restrict__ table, float alpha, float4 *__restrict__ p4, size_t nq, size_t per_cb_q) {
    extern __shared__ float smem[];
    float *x_s = smem, *tab_s = smem + CBT;
    if (threadIdx.x < 256) tab_s[threadIdx.x] = table[threadIdx.x] * alpha;
    const size_t step = (size_t)NT * U;
    size_t q = (nq * blockIdx.x / gridDim.x) / step * step;
    const size_t q_end = blockIdx.x + 1 == gridDim.x ? nq : (nq * (blockIdx.x + 1) / gridDim.x) / step * step;
    while (q < q_end) {
        const size_t cb = q / per_cb_q;
        const size_t stop = (cb + 1) * per_cb_q < q_end ? (cb + 1) * per_cb_q : q_end;
        __syncthreads();                                        // everyone is done with the previous tile
        const float4 *xg = reinterpret_cast<const float4 *>(x + cb * CBT);
        for (int i = threadIdx.x; i < CBT / 4; i += NT) reinterpret_cast<float4 *>(x_s)[i] = xg[i];
        __syncthreads();
        for (size_t qq = q + threadIdx.x; qq < stop; qq += step) {
            uint2 lc[U];
            uint32_t cd[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (qq + u * NT < stop) { lc[u] = __ldcs(lcol4 + qq + u * NT); cd[u] = __ldcs(code4 + qq + u * NT); }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (qq + u * NT < stop) {
                    float4 p;
                    p.x = x_s[lc[u].x & (CBT - 1)] * tab_s[cd[u] & 255u];
                    p.y = x_s[(lc[u].x >> 16) & (CBT - 1)] * tab_s[(cd[u] >> 8) & 255u];
                    p.z = x_s[lc[u].y & (CBT - 1)] * tab_s[(cd[u] >> 16) & 255u];
                    p.w = x_s[(lc[u].y >> 16) & (CBT - 1)] * tab_s[cd[u] >> 24];
                    __stcs(p4 + qq + u * NT, p);
                }
            }
        }
        q = stop;
    }
}

template <int U, int CBT, int NT>
static void run_pass1p(const char *what, int G, const uint2 *lcol4, const uint32_t *code4, const float *x, const float *table,
                       float *p, size_t nnz, int k, cudaEvent_t e0, cudaEvent_t e1) {
    const int smem = (CBT + 256) * 4;
    CK(cudaFuncSetAttribute((pass1p_kernel<U, CBT, NT>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    float sum = 0.0f;
    for (int it = -2; it < 8; ++it) {
        CK(cudaEventRecord(e0));
        pass1p_kernel<U, CBT, NT><<<G, NT, smem>>>(lcol4, code4, x, table, 1.0f, (float4 *)p, nnz / 4, nnz / 4 / (k / CBT));
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (it >= 0) sum += ms;
    }
    CK(cudaGetLastError());
    printf("pass 1 %-58s %7.1f us\n", what, sum / 8 * 1e3);
}

// pass 2 with 8 + 16 byte loads: a thread takes four neighbouring entries of the segment (segments start on multiples of
// four entries); D = segments loaded ahead.  The run-head test looks inside the quad only (enough for the synthetic data).
template <int D, int NT>
__global__ void __launch_bounds__(NT, 1024 / NT)
pass2v_kernel(const uint2 *__restrict__ lrow4, const float4 *__restrict__ p4, const int *__restrict__ seg_ptr,
              float *__restrict__ y, int RB, int n) {
    extern __shared__ float smem[];
    float *y_s = smem;
    int *sp = reinterpret_cast<int *>(smem + RB);
    const int rb = blockIdx.x;
    for (int i = threadIdx.x; i < RB; i += NT) y_s[i] = 0.0f;
    for (int i = threadIdx.x; i <= NCB; i += NT) sp[i] = seg_ptr[rb * NCB + i] >> 2;       // in quads
    __syncthreads();
    uint2 r[D];
    float4 v[D];
    auto load = [&](int s, uint2 &rr, float4 &vv) {
        const int i = sp[s] + threadIdx.x;
        rr = make_uint2(0xffffffffu, 0xffffffffu);
        if (i < sp[s + 1]) { rr = __ldcs(lrow4 + i); vv = __ldcs(p4 + i); }
    };
#pragma unroll
    for (int d = 0; d < D; ++d) load(d, r[d], v[d]);
    for (int s0 = 0; s0 < NCB; s0 += D) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const int s = s0 + d;
            const uint2 cr = r[d];
            const float4 cv = v[d];
            if (s + D < NCB) load(s + D, r[d], v[d]);
            if (cr.x != 0xffffffffu) {
                const uint32_t r0 = cr.x & 0xffffu, r1 = cr.x >> 16, r2 = cr.y & 0xffffu, r3 = cr.y >> 16;
                float a = y_s[r0] + cv.x;
                if (r1 == r0) a += cv.y; else { y_s[r0] = a; a = y_s[r1] + cv.y; }
                if (r2 == r1) a += cv.z; else { y_s[r1] = a; a = y_s[r2] + cv.z; }
                if (r3 == r2) a += cv.w; else { y_s[r2] = a; a = y_s[r3] + cv.w; }
                y_s[r3] = a;
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < RB; i += NT) {
        const size_t row = (size_t)rb * RB + i;
        if (row < (size_t)n) y[row] += y_s[i];
    }
}

template <int D, int NT>
static void run_pass2v(const char *what, const uint16_t *lrow, const float *p, const int *seg_ptr, float *y, int NRB, int RB, int n,
                       cudaEvent_t e0, cudaEvent_t e1) {
    const int smem = RB * 4 + (NCB + 1) * 4;
    CK(cudaFuncSetAttribute((pass2v_kernel<D, NT>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    float sum = 0.0f;
    for (int it = -2; it < 8; ++it) {
        CK(cudaEventRecord(e0));
        pass2v_kernel<D, NT><<<NRB, NT, smem>>>((const uint2 *)lrow, (const float4 *)p, seg_ptr, y, RB, n);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (it >= 0) sum += ms;
    }
    CK(cudaGetLastError());
    printf("pass 2 %-58s %7.1f us\n", what, sum / 8 * 1e3);
}

template <int U, int MODE, int CBT, int NT>
static void run_pass1(const char *what, const uint2 *lcol4, const uint32_t *code4, const float *x, const float *table, float *p,
                      size_t nnz, int k, cudaEvent_t e0, cudaEvent_t e1) {
    const int smem = (CBT + 256) * 4, items = (int)(nnz / CHUNK), cpc = items / (k / CBT);
    CK(cudaFuncSetAttribute((pass1_kernel<U, MODE, CBT, NT>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    float sum = 0.0f;
    for (int it = -2; it < 8; ++it) {
        CK(cudaEventRecord(e0));
        pass1_kernel<U, MODE, CBT, NT><<<items, NT, smem>>>(lcol4, code4, x, table, 1.0f, (float4 *)p, cpc);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (it >= 0) sum += ms;
    }
    CK(cudaGetLastError());
    printf("pass 1 %-58s %7.1f us\n", what, sum / 8 * 1e3);
}

template <int K, int D, int MODE, int NT>
static void run_pass2(const char *what, const uint16_t *lrow, const float *p, const int *seg_ptr, float *y, int NRB, int RB, int n,
                      cudaEvent_t e0, cudaEvent_t e1) {
    const int smem = RB * 4 + (NCB + 1) * 4;
    CK(cudaFuncSetAttribute((pass2_kernel<K, D, MODE, NT>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    float sum = 0.0f;
    for (int it = -2; it < 8; ++it) {
        CK(cudaEventRecord(e0));
        pass2_kernel<K, D, MODE, NT><<<NRB, NT, smem>>>(lrow, p, seg_ptr, y, RB, n);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (it >= 0) sum += ms;
    }
    CK(cudaGetLastError());
    printf("pass 2 %-58s %7.1f us\n", what, sum / 8 * 1e3);
}

int main() {
    const size_t nnz = (size_t)1 << 27;
    const int n = 1 << 23, k = 1 << 23;
    uint2 *lcol4; uint32_t *code4; float *x, *table, *p, *y; uint16_t *lrow; int *seg_ptr;
    CK(cudaMalloc(&lcol4, nnz * 2)); CK(cudaMalloc(&code4, nnz)); CK(cudaMalloc(&x, (size_t)k * 4));
    CK(cudaMalloc(&table, 1024)); CK(cudaMalloc(&p, nnz * 4)); CK(cudaMalloc(&y, (size_t)n * 4));
    CK(cudaMalloc(&lrow, nnz * 2)); CK(cudaMalloc(&seg_ptr, (size_t)(592 * NCB + 1) * 4));
    fill_pass1<<<148 * 8, 256>>>(lcol4, code4, nnz / 4);
    fill_f32<<<148 * 8, 256>>>(x, k);
    fill_f32<<<1, 256>>>(table, 256);
    CK(cudaMemset(y, 0, (size_t)n * 4));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));

    run_pass1<4, 0, 32768, 1024>("128 KB tile, 1024 threads, 4 loads in flight", lcol4, code4, x, table, p, nnz, k, e0, e1);
    run_pass1<8, 0, 8192, 256>("32 KB tile, 4 CTAs of 256 threads per SM, 8 loads", lcol4, code4, x, table, p, nnz, k, e0, e1);
    run_pass1p<4, 32768, 1024>("persistent, 128 KB tile, 148 CTAs of 1024 threads, 4 loads", 148, lcol4, code4, x, table, p, nnz, k, e0, e1);
    run_pass1p<8, 32768, 1024>("persistent, 128 KB tile, 148 CTAs of 1024 threads, 8 loads", 148, lcol4, code4, x, table, p, nnz, k, e0, e1);
    run_pass1p<4, 32768, 512>("persistent, 128 KB tile, 148 CTAs of 512 threads, 4 loads", 148, lcol4, code4, x, table, p, nnz, k, e0, e1);
    run_pass1p<8, 16384, 512>("persistent, 64 KB tile, 296 CTAs of 512 threads, 8 loads", 296, lcol4, code4, x, table, p, nnz, k, e0, e1);

    // segments: rb-major, nnz / nseg entries on average, +-12 % pseudo-random
    auto setup_pass2 = [&](int NRB, int cap) {
        const int nseg = NRB * NCB, RB = (n + NRB - 1) / NRB;
        std::vector<int> sp(nseg + 1);
        uint32_t st = 12345u;
        size_t pos = 0;
        const size_t avg = nnz / nseg;
        for (int s = 0; s < nseg; ++s) {
            sp[s] = (int)pos;
            st = st * 1664525u + 1013904223u;
            size_t len = (avg - avg / 8 + (st >> 8) % (avg / 4)) & ~(size_t)3;      // segments start on multiples of 4
            if (s == nseg - 1 || pos + len > nnz) len = nnz - pos;
            pos += len;
        }
        sp[nseg] = (int)pos;
        // the last segment took the remainder: move boundaries back until every segment has <= cap entries
        for (int s = nseg - 1; s >= 0 && sp[s + 1] - sp[s] > cap; --s) sp[s] = sp[s + 1] - cap;
        CK(cudaMemcpy(seg_ptr, sp.data(), (size_t)(nseg + 1) * 4, cudaMemcpyHostToDevice));
        fill_lrow<<<nseg, 256>>>(lrow, seg_ptr, nseg, RB);
        CK(cudaDeviceSynchronize());
        int maxlen = 0;
        for (int s = 0; s < nseg; ++s) maxlen = sp[s + 1] - sp[s] > maxlen ? sp[s + 1] - sp[s] : maxlen;
        printf("%d row blocks of %d rows, %d segments (max %d entries)\n", NRB, RB, nseg, maxlen);
        return RB;
    };
    int RB = setup_pass2(296, 2048);
    run_pass2v<1, 1024>("4 entries per thread (8 + 16 byte loads), 1024 threads, next segment", lrow, p, seg_ptr, y, 296, RB, n, e0, e1);
    run_pass2v<4, 512>("4 entries per thread, 512 threads, next 4 segments", lrow, p, seg_ptr, y, 296, RB, n, e0, e1);
    return 0;
}
