// spmv_dense.cu -- K-spmv (dense-code form): y = beta*y + alpha * W x for matrices that are not
// very sparse (>= 3/16 of the elements stored), where 5 bytes per non-zero of CSR cost more than one
// byte per ELEMENT and, above all, where the per-entry work of the row kernel (a scattered x gather
// and a bank-conflicting codebook lookup per entry) binds long before HBM does
// (profiles/r01_spmv_streaming.md).
//
// Layout: W = S^T as a dense n x kpad matrix of one-byte codes (kpad = k rounded up to 512; a
// code >= T -- 255 -- marks "no entry"), i.e. the very matrix CopyForm consumed
// (sparse-matrix.cc:21-99), rebuilt on the device from the CSR on the first product that wants it.
//
// Kernel: warp w of a CTA owns the 512 columns [512 w, 512 w + 512) of every row the CTA works on;
// lane l owns 16 consecutive ones, reads its 16 codes of a row with ONE 128-bit streaming load
// (512 contiguous bytes per warp and row) and keeps its 16 x values in REGISTERS for the whole
// kernel -- no gather at all.  The codebook (times alpha, kernel.cc:791) sits in shared memory 32
// times, entry c of lane l at word 32 c + l, so a lookup never meets a bank conflict.  Per element:
// byte extract, address, LDS, FFMA.  Four rows are in flight per warp; their lane sums are combined
// with a transposed butterfly (6 shuffles per 4 rows) and parked in shared memory per (row, warp);
// the warps never meet until the CTA's rows are done, then one thread per row adds the warps' sums
// in column order and applies beta (sparse-matrix.cc:149) -- a fixed order, hence deterministic.
//
// Reference semantics kept: an element that is not stored must not touch x.  0 * x would, for an x
// that is Inf or NaN; a warp whose 16 x 32 values are all finite (the normal case) takes the plain
// path, any other warp selects 0 for the x of absent entries.  Rows without entries keep their
// value (times beta), -0 included, through row_ptr like every other kernel.
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "device_matrix.h"

namespace sblas_b200 {

namespace {

constexpr int kColsPerWarp = 512, kRowsPerGroup = 4;

// dense[row][col[i]] = code[i]; one warp per row
__global__ void __launch_bounds__(256)
dense_code_scatter_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                          const uint8_t *__restrict__ code, int n, int64_t kpad, uint8_t *__restrict__ dense) {
    const int row = (int)(((size_t)blockIdx.x * 256 + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    const int s = row_ptr[row], e = row_ptr[row + 1];
    uint8_t *out = dense + (size_t)row * kpad;
    for (int i = s + lane; i < e; i += 32) out[col[i]] = code[i];
}

// One row's 16 elements of this lane.  The shared-memory address of a codebook word is formed by ONE
// byte permute: the table starts on a 64 KiB boundary of the shared window and entry c of lane l sits
// at byte 256 c + 4 l of it, so the address is `lane_addr` (= table + 4 l, whose byte 1 is zero) with
// the code spliced into byte 1.
__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

template <bool CLEAN>
__device__ __forceinline__ float row_block(const uint4 cw, const float (&x)[16], uint32_t lane_addr, int T) {
    const uint32_t w[4] = {cw.x, cw.y, cw.z, cw.w};
    float acc = 0.0f;
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const uint32_t addr = __byte_perm(w[q], lane_addr, 0x7604u + (b << 4));   // bytes: 4 l, code, table bits 16..31
            const float wt = lds_f32(addr);
            const float xv = CLEAN ? x[4 * q + b] : ((int)((addr >> 8) & 255u) < T ? x[4 * q + b] : 0.0f);
            acc = fmaf(xv, wt, acc);
        }
    return acc;
}

// blockDim.x = 32 * colwarps * rowsplit: warp (cw, rs) owns the columns [512 cw, 512 cw + 512) and, of the
// CTA's row groups (4 rows each), those with index = rs mod rowsplit.
template <bool CLEAN>
__device__ __forceinline__ void walk_rows(const uint8_t *__restrict__ base, int64_t kpad, int row0, int row1, int g0,
                                          int gstep, const float (&xr)[16], uint32_t lane_addr,
                                          int T, float *part, int part_rows, int cwarp, int lane, int pf_groups) {
    const int64_t stride = kpad * kRowsPerGroup * gstep;
    // lanes 0..15 ask L2 for the four 128-byte lines of the four rows `pf_groups` groups ahead: the loads that
    // follow find them there, so HBM latency is covered by L2 instead of by registers
    const int64_t pf_off = stride * pf_groups + (int64_t)(lane >> 2) * kpad + (lane & 3) * 128 - lane * 16;
    const int pf_rows = kRowsPerGroup * gstep * pf_groups + (lane >> 2);
    const uint8_t *p = base + (int64_t)(row0 + g0 * kRowsPerGroup) * kpad;
    const uint4 none = make_uint4(~0u, ~0u, ~0u, ~0u);
    auto load = [&](const uint8_t *q, int row) { return row < row1 ? __ldcs(reinterpret_cast<const uint4 *>(q)) : none; };
    int r = row0 + g0 * kRowsPerGroup;
    uint4 c0 = load(p, r), c1 = load(p + kpad, r + 1), c2 = load(p + 2 * kpad, r + 2), c3 = load(p + 3 * kpad, r + 3);
    for (; r < row1; r += kRowsPerGroup * gstep) {
        // a row's codes are replaced by those of the same row of the next group as soon as they are used:
        // about three 128-bit loads per lane stay in flight while the warp computes
        const uint8_t *pn = p + stride;
        const int rn = r + kRowsPerGroup * gstep;
        if (pf_groups && lane < 16 && r + pf_rows < row1) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + pf_off));
        const float a0 = row_block<CLEAN>(c0, xr, lane_addr, T); c0 = load(pn, rn);
        const float a1 = row_block<CLEAN>(c1, xr, lane_addr, T); c1 = load(pn + kpad, rn + 1);
        const float a2 = row_block<CLEAN>(c2, xr, lane_addr, T); c2 = load(pn + 2 * kpad, rn + 2);
        const float a3 = row_block<CLEAN>(c3, xr, lane_addr, T); c3 = load(pn + 3 * kpad, rn + 3);
        p = pn;
        // transposed butterfly: afterwards lane 8 j holds the warp's sum of row r + j
        const bool hi = lane & 16;
        const float s0 = hi ? a0 : a2, s1 = hi ? a1 : a3;                // what goes to the partner
        const float k0 = hi ? a2 : a0, k1 = hi ? a3 : a1;                // what stays: rows (0,1) below 16, (2,3) above
        const float b0 = k0 + __shfl_xor_sync(0xffffffffu, s0, 16), b1 = k1 + __shfl_xor_sync(0xffffffffu, s1, 16);
        const bool mid = lane & 8;
        float v = (mid ? b1 : b0) + __shfl_xor_sync(0xffffffffu, mid ? b0 : b1, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        const int j = lane >> 3;
        if ((lane & 7) == 0 && r + j < row1) part[(size_t)cwarp * part_rows + (r + j - row0)] = v;   // [warp][row]: the four lanes hit four banks
    }
}

__global__ void __launch_bounds__(1024, 1)
spmv_dense_kernel(const uint8_t *__restrict__ dense, int64_t kpad, const int32_t *__restrict__ row_ptr,
                  const float *__restrict__ table, int T, const float *__restrict__ x, float *__restrict__ y, int n, int k,
                  float alpha, float beta, int rows_per_cta, int colwarps, int pf_groups, int red_add) {
    extern __shared__ __align__(256) unsigned char smem_raw[];
    // the table (256 entries x 256 bytes, 32 lanes x 4 B of each used) on the next 64 KiB boundary of the
    // shared window, the per-(row, warp) sums behind it
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t tab_addr = (sbase + 65535u) & ~65535u;
    float *tab = reinterpret_cast<float *>(smem_raw + (tab_addr - sbase));
    float *part = tab + 256 * 64;                                       // [colwarps][rows_per_cta]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cwarp = warp % colwarps, rsplit = warp / colwarps, nsplit = (blockDim.x >> 5) / colwarps;
    const int row0 = blockIdx.x * rows_per_cta;
    const int row1 = min(n, row0 + rows_per_cta);
    // this lane's 16 x values, for good (four 128-bit loads when x allows it)
    float xr[16];
    const int c0 = cwarp * kColsPerWarp + lane * 16;
    if ((reinterpret_cast<uintptr_t>(x) & 15) == 0 && c0 + 16 <= k) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 v = __ldg(reinterpret_cast<const float4 *>(x + c0) + i);
            xr[4 * i] = v.x; xr[4 * i + 1] = v.y; xr[4 * i + 2] = v.z; xr[4 * i + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) xr[i] = (c0 + i < k) ? __ldg(x + c0 + i) : 0.0f;
    }
    bool finite = true;
#pragma unroll
    for (int i = 0; i < 16; ++i) finite = finite && (fabsf(xr[i]) <= 3.402823466e38f);   // false for Inf and NaN
    const bool clean = __all_sync(0xffffffffu, finite);
    // the values the CTA's rows update and their lengths are asked for now and parked in shared memory:
    // nothing is left to fetch when the sums are ready
    float *s_c = part + (size_t)rows_per_cta * colwarps;
    int *s_len = reinterpret_cast<int *>(s_c + rows_per_cta);
    const bool red = red_add && beta == 1.0f;              // y in a peer GPU's memory: sums added by reductions there, y never read
    for (int t = threadIdx.x; t < row1 - row0; t += blockDim.x) {
        s_len[t] = row_ptr[row0 + t + 1] - row_ptr[row0 + t];
        s_c[t] = red ? 0.0f : y[row0 + t];
    }
    for (int i = threadIdx.x; i < 256 * 32; i += blockDim.x)
        tab[(i >> 5) * 64 + (i & 31)] = ((i >> 5) < T) ? table[i >> 5] * alpha : 0.0f;
    __syncthreads();
    const uint32_t lane_addr = tab_addr + lane * 4;
    if (clean) walk_rows<true>(dense + c0, kpad, row0, row1, rsplit, nsplit, xr, lane_addr, T, part, rows_per_cta, cwarp, lane, pf_groups);
    else walk_rows<false>(dense + c0, kpad, row0, row1, rsplit, nsplit, xr, lane_addr, T, part, rows_per_cta, cwarp, lane, pf_groups);
    __syncthreads();
    for (int t = threadIdx.x; t < row1 - row0; t += blockDim.x) {
        const int len = s_len[t];
        if (len == 0 && beta == 1.0f) continue;             // column without entries, no scale: untouched
        float c = s_c[t];
        if (beta != 1.0f) c *= beta;
        float sum = 0.0f;
        for (int w = 0; w < colwarps; ++w) sum += part[(size_t)w * rows_per_cta + t];
        if (red) { if (len) asm volatile("red.global.add.f32 [%0], %1;" ::"l"(y + row0 + t), "f"(sum) : "memory"); continue; }
        y[row0 + t] = len ? c + sum : c;
    }
}

}  // namespace

// Build the dense code matrix from the CSR (once per handle, on the stream of the first product that wants it).
static int build_dense_codes(sblas_b200_matrix *mat, cudaStream_t stream);
int ensure_dense_codes(sblas_b200_matrix *mat, cudaStream_t stream) {
    if (mat->dense_state == 1) return SBLAS_B200_OK;
    if (mat->dense_state < 0) return 1;
    return build_dense_codes(mat, stream);
}

static int build_dense_codes(sblas_b200_matrix *mat, cudaStream_t stream) {
    const int64_t kpad = ((int64_t)mat->k + kColsPerWarp - 1) / kColsPerWarp * kColsPerWarp;
    const size_t bytes = (size_t)mat->n * kpad;
    uint8_t *d = nullptr;
    cudaError_t e = cudaMalloc((void **)&d, bytes);
    if (e != cudaSuccess) {                                  // no room: stay with the row kernel
        cudaGetLastError();
        mat->dense_state = -1;
        return 1;
    }
    SB_CUDA(cudaMemsetAsync(d, 0xff, bytes, stream));
    SB_LAUNCH(dense_code_scatter_kernel, (unsigned)(((size_t)mat->n * 32 + 255) / 256), 256, 0, stream, mat->d_row_ptr,
              mat->d_col, mat->d_code, mat->n, kpad, d);
    SB_CHECK_LAUNCH();
    mat->d_dense_code = d;
    mat->dense_kpad = kpad;
    mat->dense_state = 1;
    mat->format_bytes += (int64_t)bytes;
    return SBLAS_B200_OK;
}

// 0 = done, < 0 = error, 1 = not applicable (caller runs the row kernel)
int run_spmv_dense(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta, cudaStream_t stream) {
    const char *env = getenv("SBLAS_B200_SPMV_DENSE");       // 0 never, 1 whenever possible (read per call: tests flip it)
    const int mode = (env && *env) ? atoi(env) : -1;
    if (mode == 0 || mat->dense_state < 0 || !mat->csr_sorted) return 1;     // (duplicate columns would collapse)
    if (mat->k > 32 * kColsPerWarp || mat->n < 1) return 1;  // one warp per 512 columns, 32 warps per CTA
    // measured crossover with the row kernel (profiles/r01d_dense_spmv.md): 16 % of the elements stored at 16384^2,
    // about 20 % at 4096^2
    if (mode != 1 && (double)mat->nnz * 16.0 < 3.0 * (double)mat->n * (double)mat->k) return 1;
    if (mat->dense_state == 0) {
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
        if (cap != cudaStreamCaptureStatusNone) return 1;    // no allocation inside a capture
        const int rc = build_dense_codes(mat, stream);
        if (rc != SBLAS_B200_OK) return rc;
    }
    static int sms = 0;
    if (!sms && (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mat->device) != cudaSuccess || sms <= 0)) sms = 148;
    // One CTA per SM: colwarps warps side by side cover a row, rowsplit such sets share the CTA's row
    // groups (short rows would otherwise leave an SM with a few warps and no loads in flight)
    const int colwarps = (int)(mat->dense_kpad / kColsPerWarp);
    const int rowsplit = std::max(1, 32 / colwarps);
    int rows_per_cta = (int)(((int64_t)mat->n + sms - 1) / sms);
    rows_per_cta = std::max(kRowsPerGroup, (rows_per_cta + kRowsPerGroup - 1) / kRowsPerGroup * kRowsPerGroup);
    const size_t tab_bytes = 2 * 65536, max_smem = 200 * 1024;            // the table and the slack to reach a 64 KiB boundary
    const int max_rows = (int)((max_smem - tab_bytes) / (4 * (colwarps + 2)));   // shared memory for the per-(row, warp) sums, c and len
    rows_per_cta = std::min(rows_per_cta, max_rows / kRowsPerGroup * kRowsPerGroup);
    const size_t smem = tab_bytes + (size_t)rows_per_cta * (colwarps + 2) * 4;
    static bool configured_on[64] = {false};
    bool &configured = configured_on[mat->device & 63];
    if (!configured) {
        SB_CUDA(cudaFuncSetAttribute(spmv_dense_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_smem));
        configured = true;
    }
    const unsigned grid = (unsigned)((mat->n + rows_per_cta - 1) / rows_per_cta);
    static const int pf_groups = [] { const char *e = getenv("SBLAS_B200_DENSE_PF"); return (e && *e) ? atoi(e) : 2; }();   // row groups prefetched ahead into L2
    struct Timer {                                           // event pair around the dominant kernel when profiling
        sblas_b200_matrix *mat; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
        Timer(sblas_b200_matrix *m_, cudaStream_t s_) : mat(m_), st(s_) {
            if (mat->profiling) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, st); }
        }
        ~Timer() { if (a) { cudaEventRecord(b, st); mat->prof_events.push_back(std::make_pair(a, b)); } }
    } timer(mat, stream);
    note_product(mat, "spmv_dense", 1, (double)mat->n * (double)mat->dense_kpad / (double)mat->nnz, (int64_t)4 * (mat->n + 1));
    SB_LAUNCH(spmv_dense_kernel, grid, colwarps * rowsplit * 32, smem, stream, mat->d_dense_code, mat->dense_kpad, mat->d_row_ptr,
              mat->d_table, mat->T, x, y, mat->n, mat->k, alpha, beta, rows_per_cta, colwarps, pf_groups, is_peer_pointer(y) ? 1 : 0);
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200
