// product.cu -- the sparse x dense product on the device:  C = beta*C + alpha*A*S.
//
// Replaces the CPU micro-kernels the reference's AddMatMat drives
// (reference src/sparse/sparse-matrix.cc:139-194 -> src/sparse/kernel.cc):
//   sblas_beta_operation_kernel      kernel.cc:10-29    -> fused into the epilogues (scale_kernel when alpha == 0)
//   sblas_trans_kernel               kernel.cc:31-187   -> transpose_in_kernel (A -> X = A^T) + smem-tile transposed stores
//   sblas_kernel_operation_trans_ex  kernel.cc:771-800  -> done once at upload (src/utils/device_format.cu)
//   sblas_kernel_mul_naive           kernel.cc:371-584  -> spmv_* (m == 1) and spmm_* (m >= 2) below
//
// North-star notation: Y[n x m] = W X with W = S^T (CSR, 1-byte codes), X = A^T, Y = C^T.
// Arithmetic kept from the reference: the weight is fl(table[code]*alpha) (alpha folded
// first, kernel.cc:791,794); C is scaled by beta only when beta != 1 (sparse-matrix.cc:149)
// and then multiplies even for beta == 0; alpha == 0 means "scale only" (:152).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "device_matrix.h"

namespace sblas_b200 {

// ---- memory access helpers -------------------------------------------------------------
// Index/value streams are read once: streaming loads (evict-first) keep them from pushing
// the gathered x / X rows out of L2.
__device__ __forceinline__ int4 ld_stream_int4(const int4 *p) { return __ldcs(p); }
__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t *p) { return __ldcs(p); }

// x gather: through L1 (ld.global.nc) or L2-only (ld.global.cg); chosen per launch (uniform)
__device__ __forceinline__ float ld_x(const float *p, int cg) { return cg ? __ldcg(p) : __ldg(p); }

// C in a peer GPU's memory (sblas_b200_peer_open) and beta == 1: the sum is added by a reduction at the owner's memory
// (one rounding, like the add it replaces) instead of reading C across the link: the RED instances of the SpMV
// kernels (a run-time flag in the one kernel cost the local case 2.4 % on the power-law matrix).
__device__ __forceinline__ void red_add(float *p, float v) { asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory"); }

// ---- K-scale: C *= beta (alpha == 0 path; sparse-matrix.cc:149-152) ----------------------
__global__ void __launch_bounds__(256)
scale_kernel(float *__restrict__ c, int m, int n, int ldc, float beta) {
    const int j = blockIdx.x * 256 + threadIdx.x;
    if (j >= n) return;
    for (int i = blockIdx.y; i < m; i += gridDim.y) c[(size_t)i * ldc + j] *= beta;
}

// ---- K-spmv: row-split, sub-warp per row ---------------------------------------------------
// LPR lanes share one W row.  A lane reads one aligned block of four column indices
// (128-bit) and four codes (32-bit) per iteration and gathers four x values; partial
// sums are combined with __shfl_xor.  Rows longer than long_threshold are left to the
// split-row kernels below.
template <int LPR, bool RED>
__global__ void __launch_bounds__(256)
spmv_rows_kernel(const int32_t *__restrict__ row_ptr, const int4 *__restrict__ col4,
                 const uint32_t *__restrict__ code4, const float *__restrict__ table, int T,
                 const float *__restrict__ x, float *__restrict__ y, int n, float alpha, float beta,
                 int long_threshold, int xcg) {
    __shared__ float s_tab[256];
    const int sub = threadIdx.x & (LPR - 1);
    const int row = (int)(((size_t)blockIdx.x * 256 + threadIdx.x) / LPR);
    int s = 0, e = 0;
    if (row < n) { s = row_ptr[row]; e = row_ptr[row + 1]; }      // before the barrier: overlaps the table load
    s_tab[threadIdx.x] = (threadIdx.x < T) ? table[threadIdx.x] * alpha : 0.0f;    // val_table[code]*alpha
    __syncthreads();
    const int len = e - s;
    const bool mine = row < n && len <= long_threshold;
    // the value this row updates is asked for now, so that its round trip overlaps the walk
    float c0 = 0.0f;
    if (!RED && mine && sub == 0 && !(len == 0 && beta == 1.0f)) c0 = y[row];
    float acc = 0.0f;
    if (mine && len > 0) {
        const int b1 = (e - 1) >> 2;
        for (int b = (s >> 2) + sub; b <= b1; b += LPR) {
            const int4 c = ld_stream_int4(col4 + b);
            const uint32_t cd = ld_stream_u32(code4 + b);
            const int base = b << 2;
            if (base >= s && base + 3 < e) {            // interior block: four independent gathers
                const float x0 = ld_x(x + c.x, xcg), x1 = ld_x(x + c.y, xcg), x2 = ld_x(x + c.z, xcg), x3 = ld_x(x + c.w, xcg);
                acc = fmaf(x0, s_tab[cd & 255u], acc);
                acc = fmaf(x1, s_tab[(cd >> 8) & 255u], acc);
                acc = fmaf(x2, s_tab[(cd >> 16) & 255u], acc);
                acc = fmaf(x3, s_tab[cd >> 24], acc);
            } else {                                    // first / last block of the row
                if (base + 0 >= s && base + 0 < e) acc = fmaf(ld_x(x + c.x, xcg), s_tab[cd & 255u], acc);
                if (base + 1 >= s && base + 1 < e) acc = fmaf(ld_x(x + c.y, xcg), s_tab[(cd >> 8) & 255u], acc);
                if (base + 2 >= s && base + 2 < e) acc = fmaf(ld_x(x + c.z, xcg), s_tab[(cd >> 16) & 255u], acc);
                if (base + 3 >= s && base + 3 < e) acc = fmaf(ld_x(x + c.w, xcg), s_tab[cd >> 24], acc);
            }
        }
    }
#pragma unroll
    for (int o = LPR >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (mine && sub == 0) {
        if (len == 0 && beta == 1.0f) return;           // column without entries, no scale: untouched
        if (RED) { red_add(y + row, acc); return; }          // beta == 1 (run_spmv)
        if (beta != 1.0f) c0 *= beta;
        y[row] = len ? c0 + acc : c0;                   // no entry: exactly the scaled value (keeps -0)
    }
}

// Same, on the length-sorted copy of the short rows (skewed matrices): position i of the sorted
// order starts on a block boundary, so only a row's last block needs masking, and the rows of one
// warp are about equally long.  y is addressed through the order list.
template <int LPR, bool RED>
__global__ void __launch_bounds__(256)
spmv_sorted_kernel(const int32_t *__restrict__ blk_ptr, const int32_t *__restrict__ slen,
                   const int32_t *__restrict__ order, const int4 *__restrict__ col4,
                   const uint32_t *__restrict__ code4, const float *__restrict__ table, int T,
                   const float *__restrict__ x, float *__restrict__ y, int n_ordered, float alpha, float beta, int xcg) {
    __shared__ float s_tab[256];
    const int sub = threadIdx.x & (LPR - 1);
    const int i = (int)(((size_t)blockIdx.x * 256 + threadIdx.x) / LPR);
    int b0 = 0, len = 0, row = 0;
    if (i < n_ordered) { b0 = blk_ptr[i]; len = slen[i]; row = order[i]; }
    s_tab[threadIdx.x] = (threadIdx.x < T) ? table[threadIdx.x] * alpha : 0.0f;    // val_table[code]*alpha
    __syncthreads();
    float c0 = 0.0f;                                        // asked for now: the round trip overlaps the walk
    if (!RED && i < n_ordered && sub == 0 && !(len == 0 && beta == 1.0f)) c0 = y[row];
    float acc = 0.0f;
    const int nfull = len >> 2, tail = len & 3;
    for (int b = sub; b < nfull; b += LPR) {
        const int4 c = ld_stream_int4(col4 + b0 + b);
        const uint32_t cd = ld_stream_u32(code4 + b0 + b);
        const float x0 = ld_x(x + c.x, xcg), x1 = ld_x(x + c.y, xcg), x2 = ld_x(x + c.z, xcg), x3 = ld_x(x + c.w, xcg);
        acc = fmaf(x0, s_tab[cd & 255u], acc);
        acc = fmaf(x1, s_tab[(cd >> 8) & 255u], acc);
        acc = fmaf(x2, s_tab[(cd >> 16) & 255u], acc);
        acc = fmaf(x3, s_tab[cd >> 24], acc);
    }
    if (tail && (nfull % LPR) == sub) {                     // the lane whose turn the last block is
        const int4 c = ld_stream_int4(col4 + b0 + nfull);
        const uint32_t cd = ld_stream_u32(code4 + b0 + nfull);
        acc = fmaf(ld_x(x + c.x, xcg), s_tab[cd & 255u], acc);
        if (tail > 1) acc = fmaf(ld_x(x + c.y, xcg), s_tab[(cd >> 8) & 255u], acc);
        if (tail > 2) acc = fmaf(ld_x(x + c.z, xcg), s_tab[(cd >> 16) & 255u], acc);
    }
#pragma unroll
    for (int o = LPR >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (i < n_ordered && sub == 0) {
        if (len == 0 && beta == 1.0f) return;               // column without entries, no scale: untouched
        if (RED) { red_add(y + row, acc); return; }          // beta == 1 (run_spmv)
        if (beta != 1.0f) c0 *= beta;
        y[row] = len ? c0 + acc : c0;                       // no entry: exactly the scaled value (keeps -0)
    }
}

// Split-row path: one CTA per segment (<= kLongSegment non-zeros) of a long row ...
__global__ void __launch_bounds__(256)
spmv_long_segment_kernel(const int32_t *__restrict__ seg_begin, const int32_t *__restrict__ seg_end,
                         const int4 *__restrict__ col4, const uint32_t *__restrict__ code4,
                         const float *__restrict__ table, int T, const float *__restrict__ x, float alpha,
                         float *__restrict__ partial, int xcg) {
    __shared__ float s_tab[256];
    __shared__ float s_red[8];
    s_tab[threadIdx.x] = (threadIdx.x < T) ? table[threadIdx.x] * alpha : 0.0f;
    __syncthreads();
    const int s = seg_begin[blockIdx.x], e = seg_end[blockIdx.x];
    float acc = 0.0f;
    const int b1 = (e - 1) >> 2;
    for (int b = (s >> 2) + threadIdx.x; b <= b1; b += 256) {
        const int4 c = ld_stream_int4(col4 + b);
        const uint32_t cd = ld_stream_u32(code4 + b);
        const int base = b << 2;
        if (base + 0 >= s && base + 0 < e) acc = fmaf(ld_x(x + c.x, xcg), s_tab[cd & 255u], acc);
        if (base + 1 >= s && base + 1 < e) acc = fmaf(ld_x(x + c.y, xcg), s_tab[(cd >> 8) & 255u], acc);
        if (base + 2 >= s && base + 2 < e) acc = fmaf(ld_x(x + c.z, xcg), s_tab[(cd >> 16) & 255u], acc);
        if (base + 3 >= s && base + 3 < e) acc = fmaf(ld_x(x + c.w, xcg), s_tab[cd >> 24], acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += s_red[w];
        partial[blockIdx.x] = t;
    }
}

// ... and one thread per long row adds its segment sums in segment order (deterministic).
__global__ void __launch_bounds__(128)
spmv_long_finish_kernel(const int32_t *__restrict__ long_rows, const int32_t *__restrict__ long_seg_ptr,
                        const float *__restrict__ partial, float *__restrict__ y, float beta, int n_long, int red) {
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= n_long) return;
    float t = 0.0f;
    for (int s = long_seg_ptr[i]; s < long_seg_ptr[i + 1]; ++s) t += partial[s];
    const int row = long_rows[i];
    if (red) { red_add(y + row, t); return; }               // beta == 1 (run_spmv)
    float c0 = y[row];
    if (beta != 1.0f) c0 *= beta;
    y[row] = c0 + t;
}

// ---- K-transpose-in: X[k][mp] = A^T, zero padded to mp (multiple of 4) ---------------------
// The function of the reference's first sblas_trans_kernel call (sparse-matrix.cc:181).
__global__ void __launch_bounds__(256)
transpose_in_kernel(const float *__restrict__ a, int m, int k, int lda, float *__restrict__ xt, int mp) {
    __shared__ float tile[32][33];
    const int r0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int ii = ty; ii < 32; ii += 8) {
        const int i = i0 + ii, r = r0 + tx;
        tile[ii][tx] = (i < m && r < k) ? a[(size_t)i * lda + r] : 0.0f;
    }
    __syncthreads();
    for (int rr = ty; rr < 32; rr += 8) {
        const int r = r0 + rr, i = i0 + tx;
        if (r < k && i < mp) xt[(size_t)r * mp + i] = tile[tx][rr];
    }
}

// ---- K-spmm (gather form): warp per W row, lanes over the dense columns --------------------
// LPN lanes cover one non-zero's X row segment as float4 (tile of MT = 4*LPN dense
// columns); a warp therefore works on 32/LPN non-zeros of its row at once ("slots"),
// reduced with __shfl_xor at the end of the row.  X rows are read straight from L2/L1.
// The 32 x MT result tile goes through shared memory so that C (row-major m x n, i.e.
// Y^T) is written as 128-byte rows along j -- the function of the reference's two
// per-panel C transposes (sparse-matrix.cc:184,189).
template <int LPN>
__global__ void __launch_bounds__(256)
spmm_gather_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                   const uint8_t *__restrict__ code, const float *__restrict__ table, int T,
                   const float *__restrict__ xt, int mp, float *__restrict__ c, int ldc, int n, int m,
                   float alpha, float beta) {
    constexpr int MT = 4 * LPN, SLOTS = 32 / LPN;
    __shared__ float s_tab[256];
    __shared__ float s_tile[32][MT + 1];
    __shared__ int s_len[32];
    s_tab[threadIdx.x] = (threadIdx.x < T) ? table[threadIdx.x] * alpha : 0.0f;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = lane / LPN, chunk = lane % LPN;
    const int m0 = blockIdx.y * MT, j0 = blockIdx.x * 32;
    const bool chunk_ok = m0 + chunk * 4 < mp;
    const float *xbase = xt + m0 + chunk * 4;
#pragma unroll 1
    for (int rr = 0; rr < 4; ++rr) {
        const int jl = warp * 4 + rr, j = j0 + jl;
        int s = 0, e = 0;
        if (j < n) { s = row_ptr[j]; e = row_ptr[j + 1]; }
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int base = s; base < e; base += 32) {
            const int idx = base + lane;
            int cc = 0;
            float w = 0.0f;
            if (idx < e) { cc = col[idx]; w = s_tab[code[idx]]; }
            const int cnt = min(32, e - base);
            // U independent gathers in flight per lane before the first FMA needs its operand
            constexpr int U = 4;
            for (int t = 0; t < cnt; t += SLOTS * U) {
                float4 xv[U];
                float wv[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int src = t + u * SLOTS + slot;           // entry of this slot in step u
                    const int r = __shfl_sync(0xffffffffu, cc, src & 31);
                    wv[u] = __shfl_sync(0xffffffffu, w, src & 31);
                    xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (chunk_ok && src < cnt) xv[u] = __ldg(reinterpret_cast<const float4 *>(xbase + (size_t)r * mp));
                    else wv[u] = 0.0f;
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    acc.x = fmaf(wv[u], xv[u].x, acc.x);
                    acc.y = fmaf(wv[u], xv[u].y, acc.y);
                    acc.z = fmaf(wv[u], xv[u].z, acc.z);
                    acc.w = fmaf(wv[u], xv[u].w, acc.w);
                }
            }
        }
#pragma unroll
        for (int o = LPN; o < 32; o <<= 1) {
            acc.x += __shfl_xor_sync(0xffffffffu, acc.x, o);
            acc.y += __shfl_xor_sync(0xffffffffu, acc.y, o);
            acc.z += __shfl_xor_sync(0xffffffffu, acc.z, o);
            acc.w += __shfl_xor_sync(0xffffffffu, acc.w, o);
        }
        if (slot == 0) {
            s_tile[jl][chunk * 4 + 0] = acc.x;
            s_tile[jl][chunk * 4 + 1] = acc.y;
            s_tile[jl][chunk * 4 + 2] = acc.z;
            s_tile[jl][chunk * 4 + 3] = acc.w;
        }
        if (lane == 0) s_len[jl] = e - s;
    }
    __syncthreads();
    const int jl = threadIdx.x & 31, j = j0 + jl;
    if (j >= n) return;
    if (s_len[jl] == 0 && beta == 1.0f) return;         // S column without entries and no scale: untouched
    for (int i = threadIdx.x >> 5; i < MT; i += 8) {
        const int gi = m0 + i;
        if (gi >= m) break;
        float *p = c + (size_t)gi * ldc + j;
        float c0 = *p;
        if (beta != 1.0f) c0 *= beta;
        *p = s_len[jl] ? c0 + s_tile[jl][i] : c0;        // no entry: exactly the scaled value (keeps -0)
    }
}

// ---------------------------------------------------------------------------------------------
// dispatch
// ---------------------------------------------------------------------------------------------
// Brackets the dominant kernel with an event pair while the handle is being profiled.
class KernelTimer {
public:
    KernelTimer(sblas_b200_matrix *mat, cudaStream_t stream) : mat_(mat), stream_(stream), on_(mat->profiling) {
        if (!on_) return;
        cudaEventCreate(&a_);
        cudaEventCreate(&b_);
        cudaEventRecord(a_, stream_);
    }
    ~KernelTimer() {
        if (!on_) return;
        cudaEventRecord(b_, stream_);
        mat_->prof_events.push_back(std::make_pair(a_, b_));
    }
private:
    sblas_b200_matrix *mat_;
    cudaStream_t stream_;
    bool on_;
    cudaEvent_t a_ = nullptr, b_ = nullptr;
};

// x gathers through L1 (ld.global.nc) when x fits there, L2-only (ld.global.cg) otherwise:
// measured 3 % faster on the 2^23-column power-law case.  SBLAS_B200_XLOAD = c | n overrides.
static int x_load_mode(const sblas_b200_matrix *mat) {
    static int forced = -2;
    if (forced == -2) {
        const char *env = getenv("SBLAS_B200_XLOAD");
        forced = !env ? -1 : (env[0] == 'c' ? 1 : 0);
    }
    if (forced >= 0) return forced;
    return mat->k > 32768 ? 1 : 0;
}

template <int LPR, bool RED>
static void launch_spmv_rows(const sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta, int xcg,
                             cudaStream_t stream) {
    char name[64];
    snprintf(name, sizeof(name), "%s<%d>%s", mat->d_row_order ? "spmv_sorted" : "spmv_rows", LPR, RED ? " red" : "");
    note_product(const_cast<sblas_b200_matrix *>(mat), name, 1, 5.0, mat->d_row_order ? (int64_t)12 * mat->n : (int64_t)4 * (mat->n + 1));
    if (mat->d_row_order) {                                 // skewed row lengths: the sorted copy
        const size_t threads = (size_t)mat->n_ordered * LPR;
        SB_LAUNCH((spmv_sorted_kernel<LPR, RED>), (unsigned)((threads + 255) / 256), 256, 0, stream, mat->d_sblk_ptr, mat->d_slen,
                  mat->d_row_order, mat->d_scol, mat->d_scode, mat->d_table, mat->T, x, y, mat->n_ordered, alpha, beta,
                  xcg);
        return;
    }
    const size_t threads = (size_t)mat->n * LPR;
    const unsigned grid = (unsigned)((threads + 255) / 256);
    SB_LAUNCH((spmv_rows_kernel<LPR, RED>), grid, 256, 0, stream, mat->d_row_ptr, reinterpret_cast<const int4 *>(mat->d_col),
              reinterpret_cast<const uint32_t *>(mat->d_code), mat->d_table, mat->T, x, y, mat->n, alpha, beta,
              mat->long_threshold, xcg);
}

static int run_spmv_sorted_to_peer(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta,
                                   cudaStream_t stream);

static int run_spmv(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta, cudaStream_t stream,
                    bool staged = false) {
    // matrices with >= 3/16 of their elements stored: one byte per element, x in registers (spmv_dense.cu)
    const int drc = staged ? 1 : run_spmv_dense(mat, x, y, alpha, beta, stream);
    if (drc <= 0) return drc;
    // large matrices in between: straight from the reference's own byte streams, x read front to back (spmv_native.cu)
    if (!staged) {
        const int nrc = run_spmv_native(mat, x, y, alpha, beta, stream);
        if (nrc <= 0) return nrc;
    }
    // the length-sorted copy visits rows in sorted order: into peer memory its row sums would travel as scattered
    // 4-byte packets (8 GPUs into one: 3.3 ms instead of 0.6) -- they are formed locally and added out in row order
    if (!staged && mat->d_row_order && is_peer_pointer(y)) return run_spmv_sorted_to_peer(mat, x, y, alpha, beta, stream);
    const int xcg = x_load_mode(mat);
    const bool red = beta == 1.0f && is_peer_pointer(y);    // y across NVLink: add at the owner, do not read it
    // Long rows: their segment sums only read x, so they are formed on a side stream beside the kernel(s) of the other rows
    // (config 4: 25 us of 437 when they ran behind them) and joined before the finish kernel adds them into y.
    // SBLAS_B200_LONG_OVERLAP=0 keeps everything on one stream.
    bool forked = false;
    if (mat->n_long) {
        int rc = ensure_buffer(&mat->d_partial, &mat->partial_bytes, (size_t)mat->n_long_seg * 4);
        if (rc != SBLAS_B200_OK) return rc;
        static const bool overlap = [] { const char *e = getenv("SBLAS_B200_LONG_OVERLAP"); return !(e && *e == '0'); }();
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
        if (overlap && !mat->side_stream && cap == cudaStreamCaptureStatusNone) {     // nothing is created inside a capture
            if (cudaStreamCreateWithFlags(&mat->side_stream, cudaStreamNonBlocking) != cudaSuccess ||
                cudaEventCreateWithFlags(&mat->side_fork, cudaEventDisableTiming) != cudaSuccess ||
                cudaEventCreateWithFlags(&mat->side_join, cudaEventDisableTiming) != cudaSuccess) {
                cudaGetLastError();
                if (mat->side_stream) { cudaStreamDestroy(mat->side_stream); mat->side_stream = nullptr; }
            }
        }
        if (overlap && mat->side_stream && mat->side_fork && mat->side_join && !mat->profiling) {
            SB_CUDA(cudaEventRecord(mat->side_fork, stream));
            SB_CUDA(cudaStreamWaitEvent(mat->side_stream, mat->side_fork, 0));
            SB_LAUNCH(spmv_long_segment_kernel, mat->n_long_seg, 256, 0, mat->side_stream, mat->d_seg_begin, mat->d_seg_end,
                      reinterpret_cast<const int4 *>(mat->d_col), reinterpret_cast<const uint32_t *>(mat->d_code),
                      mat->d_table, mat->T, x, alpha, mat->d_partial, xcg);
            SB_CUDA(cudaEventRecord(mat->side_join, mat->side_stream));
            forked = true;
        }
    }
    // huge x, scattered columns: the column-blocked copy in two streaming passes (spmv_blocked.cu); it applies beta to
    // every row, so the long rows below only add their sums
    const bool blocked = !staged && mat->blocked.state == 1;
    if (blocked) {
        const int brc = run_spmv_blocked(mat, x, y, alpha, beta, red, stream);
        if (brc != SBLAS_B200_OK) return brc;
        beta = 1.0f;
    } else {
    KernelTimer timer(mat, stream);
#define SB_SPMV_CASE(L) \
    case L: red ? launch_spmv_rows<L, true>(mat, x, y, alpha, beta, xcg, stream) \
                : launch_spmv_rows<L, false>(mat, x, y, alpha, beta, xcg, stream); break
    switch (mat->spmv_lanes) {
    SB_SPMV_CASE(1);
    SB_SPMV_CASE(2);
    SB_SPMV_CASE(4);
    SB_SPMV_CASE(8);
    SB_SPMV_CASE(16);
    default: red ? launch_spmv_rows<32, true>(mat, x, y, alpha, beta, xcg, stream)
                 : launch_spmv_rows<32, false>(mat, x, y, alpha, beta, xcg, stream); break;
    }
#undef SB_SPMV_CASE
    }
    SB_CHECK_LAUNCH();
    if (mat->n_long) {
        if (forked) SB_CUDA(cudaStreamWaitEvent(stream, mat->side_join, 0));
        else
            SB_LAUNCH(spmv_long_segment_kernel, mat->n_long_seg, 256, 0, stream, mat->d_seg_begin, mat->d_seg_end,
                      reinterpret_cast<const int4 *>(mat->d_col), reinterpret_cast<const uint32_t *>(mat->d_code),
                      mat->d_table, mat->T, x, alpha, mat->d_partial, xcg);
        SB_LAUNCH(spmv_long_finish_kernel, (mat->n_long + 127) / 128, 128, 0, stream, mat->d_long_rows,
                  mat->d_long_seg_ptr, mat->d_partial, y, beta, mat->n_long, red ? 1 : 0);
        SB_CHECK_LAUNCH();
    }
    return SBLAS_B200_OK;
}

template <int LPN>
static void launch_spmm_gather(const sblas_b200_matrix *mat, const float *xt, int mp, float *c, int ldc, int m,
                               float alpha, float beta, cudaStream_t stream) {
    const dim3 grid((mat->n + 31) / 32, (m + 4 * LPN - 1) / (4 * LPN));
    char name[64];
    snprintf(name, sizeof(name), "spmm_gather<%d>", LPN);
    note_product(const_cast<sblas_b200_matrix *>(mat), name, m, 5.0, (int64_t)4 * (mat->n + 1));
    SB_LAUNCH(spmm_gather_kernel<LPN>, grid, 256, 0, stream, mat->d_row_ptr, mat->d_col, mat->d_code, mat->d_table,
              mat->T, xt, mp, c, ldc, mat->n, m, alpha, beta);
}

static int run_spmm(sblas_b200_matrix *mat, const float *d_a, int m, int lda, float *d_c, int ldc, float alpha,
                    float beta, cudaStream_t stream) {
    // dense enough and m >= 32: the tensor cores (src/sparse/spmm_tc.cu)
    {
        const int crc = run_spmm_tc(mat, d_a, m, lda, d_c, ldc, alpha, beta, stream);
        if (crc <= 0) return crc;                       // done (0) or failed (< 0); 1 = not for this product
    }
    // large m: X staged through shared memory (src/sparse/spmm_tile.cu) once the tile format exists
    // Small m too, unless the (row, chunk) pieces are nearly empty: measured (profiles/r01c_small_m.md) the tile
    // kernel beats the gather kernel 2-3x at m = 2..16 from ~3 entries per piece, and loses 30 % below that.
    const double per_piece = (double)mat->nnz / ((double)std::max(1, mat->n) * ((mat->k + kTileK - 1) / kTileK));
    int tile_min_m = per_piece >= 3.0 ? 2 : 17;
    if (const char *env = getenv("SBLAS_B200_TILE_MIN_M")) tile_min_m = atoi(env);
    if (m >= tile_min_m) {
        const int trc = run_spmm_tile(mat, d_a, m, lda, d_c, ldc, alpha, beta, stream);
        if (trc <= 0) return trc;                       // done (0) or failed (< 0); 1 = no tile layout, use the gather kernel
    }
    const int mp = (m + 3) & ~3;
    int rc = ensure_buffer(&mat->d_ws, &mat->ws_bytes, (size_t)mat->k * mp * 4);
    if (rc != SBLAS_B200_OK) return rc;
    const dim3 tgrid((mat->k + 31) / 32, (mp + 31) / 32);
    SB_LAUNCH(transpose_in_kernel, tgrid, 256, 0, stream, d_a, m, mat->k, lda, mat->d_ws, mp);
    KernelTimer timer(mat, stream);
    if (mp <= 4) launch_spmm_gather<1>(mat, mat->d_ws, mp, d_c, ldc, m, alpha, beta, stream);
    else if (mp <= 8) launch_spmm_gather<2>(mat, mat->d_ws, mp, d_c, ldc, m, alpha, beta, stream);
    else if (mp <= 16) launch_spmm_gather<4>(mat, mat->d_ws, mp, d_c, ldc, m, alpha, beta, stream);
    else if (mp <= 32) launch_spmm_gather<8>(mat, mat->d_ws, mp, d_c, ldc, m, alpha, beta, stream);
    else if (mp <= 64) launch_spmm_gather<16>(mat, mat->d_ws, mp, d_c, ldc, m, alpha, beta, stream);
    else launch_spmm_gather<32>(mat, mat->d_ws, mp, d_c, ldc, m, alpha, beta, stream);
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

int product_on_stream(sblas_b200_matrix *mat, const float *d_a, int32_t m, int32_t lda, float *d_c, int32_t ldc,
                      float alpha, float beta, cudaStream_t stream) {
    const int n = mat->n;
    if (m == 0 || n == 0) return SBLAS_B200_OK;
    if (alpha == 0.0f || mat->nnz == 0) {               // sparse-matrix.cc:149-152: only the beta scale remains
        if (beta != 1.0f) {
            const dim3 grid((n + 255) / 256, std::min(m, 1024));
            SB_LAUNCH(scale_kernel, grid, 256, 0, stream, d_c, m, n, ldc, beta);
            SB_CHECK_LAUNCH();
        }
        return SBLAS_B200_OK;
    }
    // Xt staging and split-row partials belong to the handle: order this product after the previous one
    // when that ran on a different stream.  Not while `stream` is being captured into a CUDA graph: a
    // capture may neither wait for work outside it nor lend its events to streams outside it; products
    // inside one capture are ordered by the capture itself, and keeping replays of the graph apart from
    // other products of the same handle is then the caller's business (include/sblas_b200.h).
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
    const bool capturing = cap != cudaStreamCaptureStatusNone;
    if (!capturing) {
        if (!mat->ws_event) SB_CUDA(cudaEventCreateWithFlags(&mat->ws_event, cudaEventDisableTiming));
        if (mat->ws_used && mat->ws_stream != stream) SB_CUDA(cudaStreamWaitEvent(stream, mat->ws_event, 0));
    }
    const int rc = m == 1 ? run_spmv(mat, d_a, d_c, alpha, beta, stream) : run_spmm(mat, d_a, m, lda, d_c, ldc, alpha, beta, stream);
    if (rc == SBLAS_B200_OK && !capturing) {
        SB_CUDA(cudaEventRecord(mat->ws_event, stream));
        mat->ws_stream = stream;
        mat->ws_used = true;
    }
    return rc;
}

static int check_product_args(const sblas_b200_matrix *mat, const float *a, int32_t m, int32_t lda, const float *c,
                              int32_t ldc) {
    if (!mat) return fail(SBLAS_B200_EINVAL, "matrix handle is null");
    if (m < 0) return fail(SBLAS_B200_EINVAL, "m = %d is negative", m);
    if (m == 0 || mat->n == 0) return SBLAS_B200_OK;
    if (!c) return fail(SBLAS_B200_EINVAL, "c is null");
    if (!a && mat->k) return fail(SBLAS_B200_EINVAL, "a is null");
    // a single dense row never uses its leading dimension (the reference ignores lda / ldc for m == 1)
    if (m > 1 && ldc < mat->n) return fail(SBLAS_B200_EINVAL, "ldc %d < n %d", ldc, mat->n);
    if (m > 1 && lda < mat->k) return fail(SBLAS_B200_EINVAL, "lda %d < k %d", lda, mat->k);
    return SBLAS_B200_OK;
}

// ---- host entry point with pinned caller buffers: copies overlapped with the product --------
// The product needs A at its start but C only in its last step (out = scale(c) + sum, every
// kernel above).  So the sum is formed in a device buffer P while C is still travelling in on a
// second stream, and one kernel then applies  out = len ? scale(c) + p : scale(c)  -- the same
// two roundings as the fused epilogues -- writing straight into the caller's (mapped) C.
// P is kept at -0.0f between calls: with beta = 1 the product kernels leave (-0) + sum = sum in it,
// bit for bit (also when the sum is -0), and the combine kernel puts -0 back as it reads.
__global__ void __launch_bounds__(256)
fill_negzero_kernel(float4 *__restrict__ p, size_t n4) {
    const float z = __int_as_float(0x80000000);
    for (size_t i = blockIdx.x * (size_t)256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256)
        p[i] = make_float4(z, z, z, z);
}

template <bool VEC>
__global__ void __launch_bounds__(256)
combine_out_kernel(const float *c_in, float *__restrict__ p, const int32_t *__restrict__ row_ptr,
                   float *out, int ldo, int m, int n, float beta) {       // c_in == out is allowed (in place)
    const float z = __int_as_float(0x80000000);
    if (VEC) {
        const int n4 = n >> 2;
        const size_t total = (size_t)m * n4;
        for (size_t idx = blockIdx.x * (size_t)256 + threadIdx.x; idx < total; idx += (size_t)gridDim.x * 256) {
            const int i = (int)(idx / n4), j = (int)(idx % n4) << 2;
            float4 cv = *reinterpret_cast<const float4 *>(c_in + (size_t)i * n + j);
            const float4 pv = *reinterpret_cast<const float4 *>(p + (size_t)i * n + j);
            *reinterpret_cast<float4 *>(p + (size_t)i * n + j) = make_float4(z, z, z, z);
            const int r0 = row_ptr[j], r1 = row_ptr[j + 1], r2 = row_ptr[j + 2], r3 = row_ptr[j + 3], r4 = row_ptr[j + 4];
            if (beta != 1.0f) { cv.x *= beta; cv.y *= beta; cv.z *= beta; cv.w *= beta; }
            if (r1 != r0) cv.x += pv.x;
            if (r2 != r1) cv.y += pv.y;
            if (r3 != r2) cv.z += pv.z;
            if (r4 != r3) cv.w += pv.w;
            *reinterpret_cast<float4 *>(out + (size_t)i * ldo + j) = cv;
        }
    } else {
        const size_t total = (size_t)m * n;
        for (size_t idx = blockIdx.x * (size_t)256 + threadIdx.x; idx < total; idx += (size_t)gridDim.x * 256) {
            const int i = (int)(idx / n), j = (int)(idx % n);
            float cv = c_in[idx];
            const float pv = p[idx];
            p[idx] = z;
            if (beta != 1.0f) cv *= beta;
            if (row_ptr[j + 1] != row_ptr[j]) cv += pv;
            out[(size_t)i * ldo + j] = cv;
        }
    }
}

// ---- SpMV of the length-sorted copy with y in a peer GPU's memory -----------------------------
// out pass for beta == 1: y[j] += p[j] for the rows that have entries, as reductions at the owner's memory -- four rows
// per reduction where all four have entries (one 512-byte request per warp), single ones otherwise; p goes back to -0.
__global__ void __launch_bounds__(256)
add_out_red_kernel(float *__restrict__ p, const int32_t *__restrict__ row_ptr, float *__restrict__ y, int n, int vec) {
    const float z = __int_as_float(0x80000000);
    const int n4 = n >> 2;
    for (int q = blockIdx.x * 256 + threadIdx.x; q < n4; q += gridDim.x * 256) {
        const int j = q << 2;
        const float4 pv = *reinterpret_cast<const float4 *>(p + j);
        *reinterpret_cast<float4 *>(p + j) = make_float4(z, z, z, z);
        const int r0 = row_ptr[j], r1 = row_ptr[j + 1], r2 = row_ptr[j + 2], r3 = row_ptr[j + 3], r4 = row_ptr[j + 4];
        if (vec && r1 != r0 && r2 != r1 && r3 != r2 && r4 != r3) {
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(y + j), "f"(pv.x), "f"(pv.y), "f"(pv.z),
                         "f"(pv.w) : "memory");
        } else {
            if (r1 != r0) red_add(y + j, pv.x);
            if (r2 != r1) red_add(y + j + 1, pv.y);
            if (r3 != r2) red_add(y + j + 2, pv.z);
            if (r4 != r3) red_add(y + j + 3, pv.w);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const int j = (n4 << 2) + threadIdx.x;
        const float pv = p[j];
        p[j] = z;
        if (row_ptr[j + 1] != row_ptr[j]) red_add(y + j, pv);
    }
}

static int run_spmv_sorted_to_peer(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta,
                                   cudaStream_t stream) {
    const int n = mat->n;
    const size_t padded = ((size_t)n + 3) & ~(size_t)3;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    SB_CUDA(cudaStreamIsCapturing(stream, &cap));
    if (cap != cudaStreamCaptureStatusNone && mat->peer_y_bytes < padded * 4)
        return fail(SBLAS_B200_EINVAL, "peer output during stream capture: call once outside the capture first");
    const size_t had = mat->peer_y_bytes;
    int rc = ensure_buffer(&mat->d_peer_y, &mat->peer_y_bytes, padded * 4);
    if (rc != SBLAS_B200_OK) return rc;
    if (mat->peer_y_bytes != had) mat->peer_y_clean = false;
    if (!mat->peer_y_clean) {
        const size_t n4 = mat->peer_y_bytes / 16;
        SB_LAUNCH(fill_negzero_kernel, (unsigned)std::min<size_t>((n4 + 255) / 256, 148 * 8), 256, 0, stream,
                  reinterpret_cast<float4 *>(mat->d_peer_y), n4);
        SB_CHECK_LAUNCH();
    }
    mat->peer_y_clean = false;                              // until the out pass has put -0 back
    rc = run_spmv(mat, x, mat->d_peer_y, alpha, 1.0f, stream, true);     // (-0) + sum = sum, bit for bit
    if (rc != SBLAS_B200_OK) return rc;
    const bool aligned = (reinterpret_cast<uintptr_t>(y) & 15) == 0;
    if (beta == 1.0f) {
        const unsigned grid = (unsigned)std::min<size_t>((padded / 4 + 255) / 256, 148 * 16);
        SB_LAUNCH(add_out_red_kernel, std::max(1u, grid), 256, 0, stream, mat->d_peer_y, mat->d_row_ptr, y, n, aligned ? 1 : 0);
    } else {                                               // y read over the link once, in whole rows of the warp
        const bool vec = (n & 3) == 0 && aligned;
        const size_t work = vec ? (size_t)n / 4 : (size_t)n;
        const unsigned grid = (unsigned)std::min<size_t>((work + 255) / 256, 148 * 16);
        if (vec) SB_LAUNCH(combine_out_kernel<true>, grid, 256, 0, stream, y, mat->d_peer_y, mat->d_row_ptr, y, n, 1, n, beta);
        else SB_LAUNCH(combine_out_kernel<false>, grid, 256, 0, stream, y, mat->d_peer_y, mat->d_row_ptr, y, n, 1, n, beta);
    }
    SB_CHECK_LAUNCH();
    mat->peer_y_clean = true;
    return SBLAS_B200_OK;
}

// A strided copy whose rows touch is one plain copy: a single DMA descriptor instead of one per row.
static cudaError_t copy_rows(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t height,
                             cudaMemcpyKind kind, cudaStream_t stream) {
    static const bool force_2d = [] { const char *e = getenv("SBLAS_B200_COPY2D"); return e && *e == '1'; }();   // A/B switch
    if (dpitch == width && spitch == width && !force_2d) return cudaMemcpyAsync(dst, src, width * height, kind, stream);
    return cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height, kind, stream);
}

// device alias of a pinned + mapped host pointer, or null (pageable memory, or SBLAS_B200_HOST_OVERLAP=0)
static float *mapped_pointer(const void *host) {
    static const int enabled = [] { const char *e = getenv("SBLAS_B200_HOST_OVERLAP"); return e ? atoi(e) : 1; }();
    if (!enabled) return nullptr;
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (attr.type != cudaMemoryTypeHost || !attr.devicePointer) return nullptr;
    return static_cast<float *>(attr.devicePointer);
}

// Rows of A / C per slice of the host path: large m is cut so that slice i+1 travels in and slice
// i-1 travels out while slice i is multiplied.  Slices are multiples of 64 rows (whole m-tiles of
// the SpMM kernel); SBLAS_B200_HOST_SLICE_ROWS overrides (0 = never slice).
static int host_slice_rows(const sblas_b200_matrix *mat, int m) {
    static const int env_rows = [] { const char *e = getenv("SBLAS_B200_HOST_SLICE_ROWS"); return e ? atoi(e) : -1; }();
    int rows = env_rows;
    if (rows < 0) {
        // (slices of 256 rows, whole tiles of the tensor-core form, were measured on config 3: 1865 us per call against
        //  1696 us with 64-row slices through the tile kernel -- the finer slices overlap PCIe better)
        // a slice is at least one m-tile of the SpMM kernel (64 rows) and carries enough bytes to be
        // worth an extra round of launches (>= 4 MB of A); measured on config 3 (profiles/r01_e2e.md)
        rows = 64;
        while ((size_t)rows * mat->k * 4 < ((size_t)4 << 20) && rows < m) rows *= 2;
        if (m < 2 * rows) rows = 0;
    }
    if (rows <= 0 || rows >= m) return m;
    rows = (rows + 63) & ~63;
    if ((m + rows - 1) / rows > 16) rows = (((m + 15) / 16) + 63) & ~63;
    return rows;
}

// First rows of the slices (and m at the end).  SBLAS_B200_HOST_SLICES="128,256,128" (read per call) overrides: slice
// heights in order, the last one repeated until m is covered.
static std::vector<int> host_slice_bounds(const sblas_b200_matrix *mat, int m) {
    std::vector<int> b(1, 0);
    if (const char *e = getenv("SBLAS_B200_HOST_SLICES")) {
        int last = 0;
        for (const char *q = e; *q && b.back() < m;) {
            char *end = nullptr;
            const long v = strtol(q, &end, 10);
            if (end == q) break;
            if (v > 0) { last = (int)std::min<long>(v, m - b.back()); b.push_back(b.back() + last); }
            q = *end ? end + 1 : end;
        }
        while (last > 0 && b.back() < m && b.size() < 64) b.push_back(std::min(m, b.back() + last));
        if (b.back() == m) return b;
        b.assign(1, 0);
    }
    const int rows = host_slice_rows(mat, m);
    while (b.back() < m) b.push_back(std::min(m, b.back() + rows));
    return b;
}

static int host_product_overlapped(sblas_b200_matrix *mat, const float *a, int m, int lda, float *c, float *c_mapped,
                                   int ldc, float alpha, float beta) {
    const int n = mat->n, k = mat->k;
    const size_t elems = (size_t)m * n, padded = (elems + 3) & ~(size_t)3;
    // m == 1 with a long y (config 4: 32 MB each way): there are no rows of A / C to slice, so y is cut into column ranges
    // instead -- x travels in, then the ranges of y; ONE product; then range after range is combined and copied out while
    // the later ranges are still travelling in.  SBLAS_B200_HOST_COLSLICE=0 keeps the single direct pass.
    static const bool colslice_on = [] { const char *e = getenv("SBLAS_B200_HOST_COLSLICE"); return !(e && *e == '0'); }();
    const bool colmode = colslice_on && m == 1 && (size_t)n * 4 >= ((size_t)8 << 20);
    std::vector<int> bounds;
    if (colmode) {
        const int parts = (int)std::min<size_t>(8, (size_t)n * 4 / ((size_t)4 << 20));
        bounds.push_back(0);
        for (int q = 1; q < parts; ++q) bounds.push_back((int)(((int64_t)n * q / parts) & ~(int64_t)1023));
        bounds.push_back(n);
    } else {
        bounds = host_slice_bounds(mat, m);
    }
    const int slices = (int)bounds.size() - 1;
    if (!mat->stream2) SB_CUDA(cudaStreamCreateWithFlags(&mat->stream2, cudaStreamNonBlocking));
    if (!mat->stream3) SB_CUDA(cudaStreamCreateWithFlags(&mat->stream3, cudaStreamNonBlocking));
    while ((int)mat->host_events.size() < 3 * slices) {
        cudaEvent_t ev;
        SB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        mat->host_events.push_back(ev);
    }
    const size_t had = mat->host_p_bytes;
    int rc = ensure_buffer(&mat->d_host_p, &mat->host_p_bytes, padded * 4);
    if (rc != SBLAS_B200_OK) return rc;
    if (mat->host_p_bytes != had) mat->host_p_clean = false;
    if (!mat->host_p_clean) {
        const size_t n4 = mat->host_p_bytes / 16;
        SB_LAUNCH(fill_negzero_kernel, (unsigned)std::min<size_t>((n4 + 255) / 256, 148 * 8), 256, 0, mat->stream,
                  reinterpret_cast<float4 *>(mat->d_host_p), n4);
        SB_CHECK_LAUNCH();
    }
    mat->host_p_clean = false;                          // until every combine kernel of this call has run
    // SBLAS_B200_HOST_TRACE=1: device-side timeline of this call (events on the three streams), printed to stderr
    static const bool trace = [] { const char *e = getenv("SBLAS_B200_HOST_TRACE"); return e && *e == '1'; }();
    std::vector<std::pair<const char *, cudaEvent_t> > marks;
    auto mark = [&](const char *what, cudaStream_t st) {
        if (!trace) return;
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        cudaEventRecord(ev, st);
        marks.push_back(std::make_pair(what, ev));
    };
    mark("start", mat->stream2);
    auto fail_sync = [&](int code) { cudaStreamSynchronize(mat->stream2); cudaStreamSynchronize(mat->stream3);
                                     cudaStreamSynchronize(mat->stream); return code; };
    auto copy_in = [&](bool is_a, int s) -> cudaError_t {
        if (colmode) {                                  // x once, then y columns [bounds[s], bounds[s + 1])
            cudaError_t ce = cudaSuccess;
            if (is_a) {
                if (s != 0) return cudaSuccess;
                ce = cudaMemcpyAsync(mat->d_host_a, a, (size_t)k * 4, cudaMemcpyHostToDevice, mat->stream2);
            } else {
                ce = cudaMemcpyAsync(mat->d_host_c + bounds[s], c + bounds[s], (size_t)(bounds[s + 1] - bounds[s]) * 4,
                                     cudaMemcpyHostToDevice, mat->stream2);
            }
            if (ce == cudaSuccess) ce = cudaEventRecord(mat->host_events[3 * s + (is_a ? 0 : 1)], mat->stream2);
            mark(is_a ? "x in" : "y range in", mat->stream2);
            return ce;
        }
        const int i0 = bounds[s], mi = bounds[s + 1] - i0;
        cudaError_t e = is_a
            ? copy_rows(mat->d_host_a + (size_t)i0 * k, (size_t)k * 4, a + (size_t)i0 * lda, (size_t)lda * 4,
                                (size_t)k * 4, mi, cudaMemcpyHostToDevice, mat->stream2)
            : copy_rows(mat->d_host_c + (size_t)i0 * n, (size_t)n * 4, c + (size_t)i0 * ldc, (size_t)ldc * 4,
                                (size_t)n * 4, mi, cudaMemcpyHostToDevice, mat->stream2);
        if (e == cudaSuccess) e = cudaEventRecord(mat->host_events[3 * s + (is_a ? 0 : 1)], mat->stream2);
        mark(is_a ? "A slice in" : "C slice in", mat->stream2);
        return e;
    };
    // copy engine order: A0, A1, C0, A2, C1, ... -- the next product's input before this one's C
    cudaError_t e = copy_in(true, 0);
    for (int s = 0; s < slices && e == cudaSuccess; ++s) {
        if (s + 1 < slices) e = copy_in(true, s + 1);
        if (e == cudaSuccess) e = copy_in(false, s);
    }
    if (e != cudaSuccess) return fail_sync(cuda_fail(e, "host path copies in", __FILE__, __LINE__));
    const bool direct = slices == 1;                    // one slice: the combine kernel writes the caller's C itself
    // (every slice written through the mapped pointer instead of copied out: 2160-2740 us on config 3 against 1700, r02)
    for (int s = 0; colmode && s < slices; ++s) {
        const int j0 = bounds[s], w = bounds[s + 1] - j0;
        if (s == 0) {
            e = cudaStreamWaitEvent(mat->stream, mat->host_events[0], 0);
            if (e != cudaSuccess) return fail_sync(cuda_fail(e, "host path wait x", __FILE__, __LINE__));
            rc = product_on_stream(mat, mat->d_host_a, 1, k, mat->d_host_p, n, alpha, 1.0f, mat->stream);
            if (rc != SBLAS_B200_OK) return fail_sync(rc);
            mark("product done", mat->stream);
        }
        e = cudaStreamWaitEvent(mat->stream, mat->host_events[3 * s + 1], 0);
        if (e != cudaSuccess) return fail_sync(cuda_fail(e, "host path wait y range", __FILE__, __LINE__));
        float *p = mat->d_host_p + j0, *cs = mat->d_host_c + j0;
        const bool vec = (w & 3) == 0 && (j0 & 3) == 0;
        const size_t work = vec ? (size_t)w / 4 : (size_t)w;
        const unsigned grid = (unsigned)std::min<size_t>((work + 255) / 256, 148 * 16);
        if (vec) SB_LAUNCH(combine_out_kernel<true>, grid, 256, 0, mat->stream, cs, p, mat->d_row_ptr + j0, cs, w, 1, w, beta);
        else SB_LAUNCH(combine_out_kernel<false>, grid, 256, 0, mat->stream, cs, p, mat->d_row_ptr + j0, cs, w, 1, w, beta);
        e = cudaGetLastError();
        mark("combine done", mat->stream);
        if (e == cudaSuccess) e = cudaEventRecord(mat->host_events[3 * s + 2], mat->stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(mat->stream3, mat->host_events[3 * s + 2], 0);
        if (e == cudaSuccess) e = cudaMemcpyAsync(c + j0, cs, (size_t)w * 4, cudaMemcpyDeviceToHost, mat->stream3);
        mark("y range out", mat->stream3);
        if (e != cudaSuccess) return fail_sync(cuda_fail(e, "host path combine / copy out", __FILE__, __LINE__));
    }
    for (int s = 0; !colmode && s < slices; ++s) {
        const int i0 = bounds[s], mi = bounds[s + 1] - i0;
        float *p = mat->d_host_p + (size_t)i0 * n, *cs = mat->d_host_c + (size_t)i0 * n;
        e = cudaStreamWaitEvent(mat->stream, mat->host_events[3 * s], 0);
        if (e != cudaSuccess) return fail_sync(cuda_fail(e, "host path wait A", __FILE__, __LINE__));
        rc = product_on_stream(mat, mat->d_host_a + (size_t)i0 * k, mi, k, p, n, alpha, 1.0f, mat->stream);
        if (rc != SBLAS_B200_OK) return fail_sync(rc);
        mark("product done", mat->stream);
        e = cudaStreamWaitEvent(mat->stream, mat->host_events[3 * s + 1], 0);
        if (e != cudaSuccess) return fail_sync(cuda_fail(e, "host path wait C", __FILE__, __LINE__));
        float *out = direct ? c_mapped + (size_t)i0 * ldc : cs;
        const int ldo = direct ? ldc : n;
        const size_t se = (size_t)mi * n;
        const bool vec = (n & 3) == 0 && (ldo & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
        const size_t work = vec ? se / 4 : se;
        const unsigned grid = (unsigned)std::min<size_t>((work + 255) / 256, 148 * 16);
        if (vec) SB_LAUNCH(combine_out_kernel<true>, grid, 256, 0, mat->stream, cs, p, mat->d_row_ptr, out, ldo, mi, n, beta);
        else SB_LAUNCH(combine_out_kernel<false>, grid, 256, 0, mat->stream, cs, p, mat->d_row_ptr, out, ldo, mi, n, beta);
        e = cudaGetLastError();
        mark("combine done", mat->stream);
        if (e == cudaSuccess && !direct) {
            e = cudaEventRecord(mat->host_events[3 * s + 2], mat->stream);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(mat->stream3, mat->host_events[3 * s + 2], 0);
            if (e == cudaSuccess)
                e = copy_rows(c + (size_t)i0 * ldc, (size_t)ldc * 4, cs, (size_t)n * 4, (size_t)n * 4, mi,
                                      cudaMemcpyDeviceToHost, mat->stream3);
            mark("slice out", mat->stream3);
        }
        if (e != cudaSuccess) return fail_sync(cuda_fail(e, "host path combine / copy out", __FILE__, __LINE__));
    }
    SB_CUDA(cudaStreamSynchronize(mat->stream));
    if (!direct) SB_CUDA(cudaStreamSynchronize(mat->stream3));
    if (trace) {
        fprintf(stderr, "host path m=%d slices=%d:", m, slices);
        for (size_t i = 1; i < marks.size(); ++i) {
            float ms = 0.0f;
            cudaEventElapsedTime(&ms, marks[0].second, marks[i].second);
            fprintf(stderr, " [%s %.1f us]", marks[i].first, ms * 1e3f);
        }
        fprintf(stderr, "\n");
        for (auto &mk : marks) cudaEventDestroy(mk.second);
    }
    mat->host_p_clean = true;
    return SBLAS_B200_OK;
}

// ---- host entry point, small SpMV with mapped caller buffers: no DMA at all ------------------
// A 16 KB copy costs 4-15 us of copy-engine latency each way, more than the product itself.  For
// m = 1 and small operands one kernel reads x through the mapped pointer (coalesced) into the
// handle's staging buffer and the SpMV kernel then updates the caller's y in place through the
// mapped pointer: two launches on one stream and one synchronisation.
__global__ void __launch_bounds__(256)
stage_in_kernel(const float *__restrict__ src, float *__restrict__ dst, int count) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < count) dst[i] = src[i];
}

static int host_product_direct(sblas_b200_matrix *mat, const float *a_mapped, float *c_mapped, float alpha, float beta) {
    SB_LAUNCH(stage_in_kernel, (mat->k + 255) / 256, 256, 0, mat->stream, a_mapped, mat->d_host_a, mat->k);
    SB_CHECK_LAUNCH();
    const int rc = product_on_stream(mat, mat->d_host_a, 1, mat->k, c_mapped, mat->n, alpha, beta, mat->stream);
    if (rc != SBLAS_B200_OK) { cudaStreamSynchronize(mat->stream); return rc; }
    SB_CUDA(cudaStreamSynchronize(mat->stream));
    return SBLAS_B200_OK;
}

// ---- product without any allocation (sblas_b200_addmatmat_dev_noalloc) ------------------------
// For callers that cannot allocate (stream capture, tight serving loops): the gather-form product above with its
// X = A^T staging in a fixed device array instead of the handle's workspace.  The array's name is a host-side
// stand-in in host code, not a device address: its address is fetched once per device with
// cudaGetSymbolAddress (that device current) and handed to the kernels.  Every call on a device shares the
// array; it keeps the previous call's contents, which is harmless: transpose_in_kernel rewrites all k x mp
// floats the gather kernel reads (padding columns as zeros) before that kernel starts, on the same stream.
static __device__ __align__(256) float g_noalloc_scratch[SBLAS_B200_NOALLOC_MAX_FLOATS];

static int noalloc_scratch(int device, float **out) {
    static std::mutex lock;
    static float *ptr_on[64] = {nullptr};
    if (device < 0 || device >= 64) return fail(SBLAS_B200_EINVAL, "device %d outside [0, 64)", device);
    std::lock_guard<std::mutex> guard(lock);
    if (!ptr_on[device]) {                                  // set-up, once per device; `device` is current (DeviceScope)
        float *p = nullptr;
        SB_CUDA(cudaGetSymbolAddress((void **)&p, g_noalloc_scratch));
        ptr_on[device] = p;
    }
    *out = ptr_on[device];
    return SBLAS_B200_OK;
}

static int product_noalloc(sblas_b200_matrix *mat, const float *d_a, int m, int lda, float *d_c, int ldc, float alpha,
                           float beta, cudaStream_t stream) {
    const int n = mat->n;
    if (m == 0 || n == 0) return SBLAS_B200_OK;
    if (alpha == 0.0f || mat->nnz == 0) {                   // sparse-matrix.cc:149-152: only the beta scale remains
        if (beta != 1.0f) {
            const dim3 grid((n + 255) / 256, std::min(m, 1024));
            SB_LAUNCH(scale_kernel, grid, 256, 0, stream, d_c, m, n, ldc, beta);
            SB_CHECK_LAUNCH();
        }
        return SBLAS_B200_OK;
    }
    const int mp = (m + 3) & ~3;
    if ((int64_t)mat->k * mp > (int64_t)SBLAS_B200_NOALLOC_MAX_FLOATS)
        return fail(SBLAS_B200_EINVAL, "sblas_b200_addmatmat_dev_noalloc: k * round_up(m, 4) = %lld floats exceeds the fixed "
                    "scratch of %lld floats (SBLAS_B200_NOALLOC_MAX_FLOATS); use sblas_b200_addmatmat_dev",
                    (long long)mat->k * mp, (long long)SBLAS_B200_NOALLOC_MAX_FLOATS);
    float *xt = nullptr;
    const int rc = noalloc_scratch(mat->device, &xt);
    if (rc != SBLAS_B200_OK) return rc;
    const dim3 tgrid((mat->k + 31) / 32, (mp + 31) / 32);
    SB_LAUNCH(transpose_in_kernel, tgrid, 256, 0, stream, d_a, m, mat->k, lda, xt, mp);
#define SB_NOALLOC_GATHER(LPN)                                                                                   \
    SB_LAUNCH(spmm_gather_kernel<LPN>, dim3((n + 31) / 32, (m + 4 * LPN - 1) / (4 * LPN)), 256, 0, stream,       \
              mat->d_row_ptr, mat->d_col, mat->d_code, mat->d_table, mat->T, xt, mp, d_c, ldc, n, m, alpha, beta)
    if (mp <= 4) SB_NOALLOC_GATHER(1);
    else if (mp <= 8) SB_NOALLOC_GATHER(2);
    else if (mp <= 16) SB_NOALLOC_GATHER(4);
    else if (mp <= 32) SB_NOALLOC_GATHER(8);
    else if (mp <= 64) SB_NOALLOC_GATHER(16);
    else SB_NOALLOC_GATHER(32);
#undef SB_NOALLOC_GATHER
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200

using namespace sblas_b200;

extern "C" {

int sblas_b200_addmatmat_dev(sblas_b200_matrix *mat, const float *d_a, int32_t m, int32_t lda, float *d_c, int32_t ldc,
                             float alpha, float beta, void *stream) {
    int rc = check_product_args(mat, d_a, m, lda, d_c, ldc);
    if (rc != SBLAS_B200_OK) return rc;
    if (m == 1) { lda = std::max(lda, mat->k); ldc = std::max(ldc, mat->n); }     // unused for one row; keeps the copies' pitches valid
    DeviceScope scope(mat->device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    std::lock_guard<std::mutex> guard(mat->lock);
    return product_on_stream(mat, d_a, m, lda, d_c, ldc, alpha, beta, static_cast<cudaStream_t>(stream));
}

int sblas_b200_addmatmat_dev_noalloc(sblas_b200_matrix *mat, const float *d_a, int32_t m, int32_t lda, float *d_c,
                                     int32_t ldc, float alpha, float beta, void *stream) {
    int rc = check_product_args(mat, d_a, m, lda, d_c, ldc);
    if (rc != SBLAS_B200_OK) return rc;
    if (m == 1) { lda = std::max(lda, mat->k); ldc = std::max(ldc, mat->n); }     // unused for one row; keeps the copies' pitches valid
    DeviceScope scope(mat->device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    return product_noalloc(mat, d_a, m, lda, d_c, ldc, alpha, beta, static_cast<cudaStream_t>(stream));
}

int sblas_b200_last_product(const sblas_b200_matrix *mat, sblas_b200_product_info *out) {
    if (!mat || !out) return fail(SBLAS_B200_EINVAL, "null argument");
    memset(out, 0, sizeof(*out));
    memcpy(out->kernel, mat->last_kernel, sizeof(out->kernel));
    out->m = mat->last_m;
    out->algorithmic_bytes = mat->last_bytes;
    out->algorithmic_flops = (int64_t)2 * mat->nnz * mat->last_m;
    out->format_bytes_per_nnz = mat->last_bytes_per_nnz;
    return SBLAS_B200_OK;
}

int sblas_b200_profile_begin(sblas_b200_matrix *mat) {
    if (!mat) return fail(SBLAS_B200_EINVAL, "matrix handle is null");
    std::lock_guard<std::mutex> guard(mat->lock);
    mat->profiling = true;
    return SBLAS_B200_OK;
}

int sblas_b200_profile_end(sblas_b200_matrix *mat, double *total_ms, int64_t *launches) {
    if (!mat) return fail(SBLAS_B200_EINVAL, "matrix handle is null");
    DeviceScope scope(mat->device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    std::lock_guard<std::mutex> guard(mat->lock);
    mat->profiling = false;
    double sum = 0.0;
    int rc = SBLAS_B200_OK;
    for (auto &ev : mat->prof_events) {
        float ms = 0.0f;
        cudaError_t e = cudaEventSynchronize(ev.second);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, ev.first, ev.second);
        if (e != cudaSuccess && rc == SBLAS_B200_OK) rc = cuda_fail(e, "profile_end", __FILE__, __LINE__);
        sum += ms;
        cudaEventDestroy(ev.first);
        cudaEventDestroy(ev.second);
    }
    if (total_ms) *total_ms = sum;
    if (launches) *launches = (int64_t)mat->prof_events.size();
    mat->prof_events.clear();
    return rc;
}

int sblas_b200_addmatmat_host(sblas_b200_matrix *mat, const float *a, int32_t m, int32_t lda, float *c, int32_t ldc,
                              float alpha, float beta) {
    int rc = check_product_args(mat, a, m, lda, c, ldc);
    if (rc != SBLAS_B200_OK) return rc;
    if (m == 1) { lda = std::max(lda, mat->k); ldc = std::max(ldc, mat->n); }     // unused for one row; keeps the copies' pitches valid
    if (m == 0 || mat->n == 0) return SBLAS_B200_OK;
    DeviceScope scope(mat->device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    std::lock_guard<std::mutex> guard(mat->lock);
    const int n = mat->n, k = mat->k;
    if (!mat->stream) SB_CUDA(cudaStreamCreateWithFlags(&mat->stream, cudaStreamNonBlocking));
    const bool need_a = alpha != 0.0f && mat->nnz != 0;
    if (beta == 1.0f && !need_a) return SBLAS_B200_OK;  // nothing changes C
    if (need_a) {
        rc = ensure_buffer(&mat->d_host_a, &mat->host_a_bytes, (size_t)m * k * 4);
        if (rc != SBLAS_B200_OK) return rc;
    }
    rc = ensure_buffer(&mat->d_host_c, &mat->host_c_bytes, (size_t)m * n * 4);
    if (rc != SBLAS_B200_OK) return rc;
    if (need_a) {                                       // pageable buffers seen before get registered (pinned_memory.cu)
        note_host_buffer(mat, a, ((size_t)(m - 1) * lda + k) * 4);
        note_host_buffer(mat, c, ((size_t)(m - 1) * ldc + n) * 4);
    }
    float *c_mapped = need_a ? mapped_pointer(c) : nullptr;
    if (c_mapped && m == 1 && (size_t)(k + n) * 4 <= ((size_t)256 << 10)) {
        static const bool direct = [] { const char *e = getenv("SBLAS_B200_HOST_DIRECT"); return !(e && *e == '0'); }();   // A/B switch
        const float *a_mapped = direct ? mapped_pointer(a) : nullptr;
        if (a_mapped) return host_product_direct(mat, a_mapped, c_mapped, alpha, beta);
    }
    if (c_mapped) return host_product_overlapped(mat, a, m, lda, c, c_mapped, ldc, alpha, beta);
    // pageable caller memory (or scale only): copies and kernels back to back on one stream
    if (need_a)
        SB_CUDA(copy_rows(mat->d_host_a, (size_t)k * 4, a, (size_t)lda * 4, (size_t)k * 4, m,
                                  cudaMemcpyHostToDevice, mat->stream));
    SB_CUDA(copy_rows(mat->d_host_c, (size_t)n * 4, c, (size_t)ldc * 4, (size_t)n * 4, m,
                              cudaMemcpyHostToDevice, mat->stream));
    rc = product_on_stream(mat, mat->d_host_a, m, k, mat->d_host_c, n, alpha, beta, mat->stream);
    if (rc != SBLAS_B200_OK) return rc;
    SB_CUDA(copy_rows(c, (size_t)ldc * 4, mat->d_host_c, (size_t)n * 4, (size_t)n * 4, m,
                              cudaMemcpyDeviceToHost, mat->stream));
    SB_CUDA(cudaStreamSynchronize(mat->stream));
    return SBLAS_B200_OK;
}

}  // extern "C"
