// copyform.cu -- SparseMatrix::CopyForm on the device.
//
// The reference builds its streams with one sequential scan of the dense code matrix
// (sparse-matrix.cc:21-99): panels of 256 S columns, inside a panel every element with a valid
// code (< valid_table_size) in order of lin = r*256 + jj, the gap to the previous valid element
// bridged by (gap-1)/255 filler entries (pos 255, code T) when it exceeds 255 (:46-51), then the
// entry itself (pos = what is left of the gap, code).  last_lin starts at 0 in every panel, and
// panels without any valid element are not recorded (:57).
//
// Here the scan is cut into chunks of 32 S rows of one panel (8192 lin slots, one CTA, one run of
// 32 consecutive slots per thread) and done in three data-parallel passes:
//   A  last valid lin of every chunk                     -> per-panel prefix max = "previous valid
//      lin" entering each chunk (0 at the panel start, as in the reference)
//   B  entries (valid elements + fillers) of every chunk -> exclusive scan = stream offset of the chunk
//   C  the same walk again, writing pos/code bytes at offset + prefix inside the chunk
// The streams stay on the device, are converted to the kernels' layout by the ordinary upload
// path (device_format.cu) without a host round trip, and are copied out once for the host object.
#include <stdlib.h>

#include <algorithm>
#include <new>
#include <vector>

#include "device_matrix.h"

struct sblas_b200_copyform_job {
    int device = 0;
    int32_t s_rows = 0, s_cols = 0, T = 0;
    int32_t npanels = 0, chunks_per_panel = 0;
    int64_t entries = 0;
    uint8_t *d_pos = nullptr, *d_val = nullptr;
    std::vector<int32_t> block_bounds, block_index_bounds;      // recorded (non-empty) panels
};

namespace sblas_b200 {

constexpr int kCfRows = 32;             // S rows per chunk
constexpr int kCfThreads = 256;         // threads per chunk; thread t owns lin slots [32 t, 32 t + 32) of the chunk
constexpr int kCfRowBytes = 256 + 4;    // padded tile row (bank spread for the per-thread runs)

__device__ __forceinline__ int block_max_256(int v, int *s_warp) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = v;
    __syncthreads();
    int r = s_warp[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) r = max(r, s_warp[w]);
    __syncthreads();
    return r;
}

// exclusive scans over the 256 threads of a CTA; *total = reduction over all threads
template <bool IS_MAX>
__device__ __forceinline__ int block_exclusive_256(int v, int *s_warp, int *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc = IS_MAX ? max(inc, t) : inc + t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    int before = 0, all = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const int t = s_warp[w];
        if (w < warp) before = IS_MAX ? max(before, t) : before + t;
        all = IS_MAX ? max(all, t) : all + t;
    }
    __syncthreads();
    int prev = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) prev = 0;
    *total = all;
    return IS_MAX ? max(before, prev) : before + prev;
}

// S(r, j) of the dense input: codes[r*stride + j] (NoTrans) or codes[j*stride + r] (Trans)
__device__ __forceinline__ void load_code_tile(const uint8_t *__restrict__ codes, int s_rows, int s_cols, int64_t stride,
                                               int trans, int r0, int j0, uint8_t (*tile)[kCfRowBytes]) {
    if (!trans) {
        const int jj = threadIdx.x, j = j0 + jj;
        for (int rr = 0; rr < kCfRows; ++rr) {
            const int r = r0 + rr;
            tile[rr][jj] = (r < s_rows && j < s_cols) ? codes[(int64_t)r * stride + j] : (uint8_t)255;
        }
    } else {
        const int rr = threadIdx.x & 31, r = r0 + rr;
        for (int jj = threadIdx.x >> 5; jj < 256; jj += kCfThreads / 32) {
            const int j = j0 + jj;
            tile[rr][jj] = (r < s_rows && j < s_cols) ? codes[(int64_t)j * stride + r] : (uint8_t)255;
        }
    }
}

// MODE 0: chunk_last[chunk] = last valid lin of the chunk (0 if none)
// MODE 1: chunk_entries[chunk] = entries the chunk emits, given chunk_carry (previous valid lin)
// MODE 2: write them at chunk_base[chunk]
template <int MODE>
__global__ void __launch_bounds__(kCfThreads)
copyform_chunk_kernel(const uint8_t *__restrict__ codes, int s_rows, int s_cols, int64_t stride, int trans, int T,
                      int chunks_per_panel, const int32_t *__restrict__ chunk_carry, int32_t *__restrict__ chunk_last,
                      int32_t *__restrict__ chunk_entries, unsigned long long *__restrict__ total_entries,
                      const int32_t *__restrict__ chunk_base, uint8_t *__restrict__ pos_out,
                      uint8_t *__restrict__ val_out) {
    __shared__ __align__(16) uint8_t tile[kCfRows][kCfRowBytes];
    __shared__ int s_warp[8];
    const int chunk = blockIdx.x;
    const int panel = chunk / chunks_per_panel, r0 = (chunk - panel * chunks_per_panel) * kCfRows, j0 = panel * 256;
    load_code_tile(codes, s_rows, s_cols, stride, trans, r0, j0, tile);
    __syncthreads();
    const int rr = threadIdx.x >> 3, jj0 = (threadIdx.x & 7) * 32;
    const int lin0 = (r0 + rr) * 256 + jj0;
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = *reinterpret_cast<const uint32_t *>(&tile[rr][jj0 + 4 * i]);
    int my_last = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const int code = (w[i >> 2] >> (8 * (i & 3))) & 255;
        if (code < T) my_last = lin0 + i;
    }
    if (MODE == 0) {
        const int last = block_max_256(my_last, s_warp);
        if (threadIdx.x == 0) chunk_last[chunk] = last;
        return;
    }
    int dummy;
    int prev = max(block_exclusive_256<true>(my_last, s_warp, &dummy), chunk_carry[chunk]);
    const int prev0 = prev;
    int count = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const int code = (w[i >> 2] >> (8 * (i & 3))) & 255;
        if (code < T) {
            const int gap = lin0 + i - prev;
            count += 1 + (gap > 255 ? (gap - 1) / 255 : 0);
            prev = lin0 + i;
        }
    }
    int total;
    const int before = block_exclusive_256<false>(count, s_warp, &total);
    if (MODE == 1) {
        if (threadIdx.x == 0) {
            chunk_entries[chunk] = total;
            atomicAdd(total_entries, (unsigned long long)total);
        }
        return;
    }
    int64_t o = (int64_t)chunk_base[chunk] + before;
    prev = prev0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        const int code = (w[i >> 2] >> (8 * (i & 3))) & 255;
        if (code < T) {
            int gap = lin0 + i - prev;
            for (; gap > 255; gap -= 255, ++o) { pos_out[o] = 255; val_out[o] = (uint8_t)T; }    // fillers (sparse-matrix.cc:46-51)
            pos_out[o] = (uint8_t)gap;
            val_out[o] = (uint8_t)code;
            ++o;
            prev = lin0 + i;
        }
    }
}

// one CTA per panel: carry[c] = max(0, last[c'] for c' < c) -- the reference's last_lin entering chunk c
__global__ void __launch_bounds__(kCfThreads)
copyform_panel_carry_kernel(const int32_t *__restrict__ chunk_last, int chunks_per_panel, int32_t *__restrict__ chunk_carry) {
    __shared__ int s_warp[8];
    const int64_t base = (int64_t)blockIdx.x * chunks_per_panel;
    int running = 0;
    for (int c0 = 0; c0 < chunks_per_panel; c0 += kCfThreads) {
        const int c = c0 + threadIdx.x;
        const int v = c < chunks_per_panel ? chunk_last[base + c] : 0;
        int all;
        const int before = block_exclusive_256<true>(v, s_warp, &all);
        if (c < chunks_per_panel) chunk_carry[base + c] = max(running, before);
        running = max(running, all);
    }
}

}  // namespace sblas_b200

using namespace sblas_b200;

namespace {
struct CfScratch {
    uint8_t *codes = nullptr;
    int32_t *last = nullptr, *carry = nullptr, *entries = nullptr, *base = nullptr;
    unsigned long long *total = nullptr;
    ~CfScratch() { cudaFree(codes); cudaFree(last); cudaFree(carry); cudaFree(entries); cudaFree(base); cudaFree(total); }
};

void destroy_job(sblas_b200_copyform_job *job) {
    if (!job) return;
    DeviceScope scope(job->device);
    cudaFree(job->d_pos);
    cudaFree(job->d_val);
    delete job;
}
}  // namespace

extern "C" {

int sblas_b200_copyform_begin(const uint8_t *codes, int32_t rows, int32_t cols, int32_t stride, int32_t T, int32_t trans,
                              int32_t device, sblas_b200_copyform_job **out, int64_t *entries, int32_t *nblocks) {
    if (!out || !entries || !nblocks) return fail(SBLAS_B200_EINVAL, "null output pointer");
    *out = nullptr; *entries = 0; *nblocks = 0;
    if (T < 0 || T > 255) return fail(SBLAS_B200_EINVAL, "valid_table_size %d outside [0,255]", T);      // sparse-matrix.cc:25
    if (rows < 0 || cols < 0 || stride < 0) return fail(SBLAS_B200_EINVAL, "negative size");
    if (stride < cols) return fail(SBLAS_B200_EINVAL, "stride %d < cols %d", stride, cols);
    if ((int64_t)rows * cols && !codes) return fail(SBLAS_B200_EINVAL, "codes is null");
    const int32_t s_rows = trans ? cols : rows, s_cols = trans ? rows : cols;
    if ((int64_t)s_rows * 256 > ((int64_t)1 << 31)) return fail(SBLAS_B200_EINVAL, "rows*256 exceeds the reference's int position range");
    DeviceScope scope(device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    sblas_b200_copyform_job *job = new (std::nothrow) sblas_b200_copyform_job();
    if (!job) return fail(SBLAS_B200_ENOMEM, "host allocation failed");
    job->device = device; job->s_rows = s_rows; job->s_cols = s_cols; job->T = T;
    *out = job;
    if (T == 0 || s_rows == 0 || s_cols == 0) return SBLAS_B200_OK;             // empty object (sparse-matrix.cc:26)
    const int npanels = (s_cols + 255) / 256, cpp = (s_rows + kCfRows - 1) / kCfRows;
    const int64_t nchunks = (int64_t)npanels * cpp;
    if (nchunks >= ((int64_t)1 << 31)) { destroy_job(job); *out = nullptr; return fail(SBLAS_B200_EINVAL, "matrix too large"); }
    job->npanels = npanels; job->chunks_per_panel = cpp;
    CfScratch sc;
    auto bail = [&](int rc) { destroy_job(job); *out = nullptr; return rc; };
    // what the reference's scan touches (sparse-matrix.cc:33-54): the last row ends after `cols` elements, not `stride`
    const size_t dense_bytes = (size_t)(rows - 1) * stride + cols;
    cudaError_t e = cudaMalloc((void **)&sc.codes, dense_bytes);
    if (e == cudaSuccess) e = cudaMemcpy(sc.codes, codes, dense_bytes, cudaMemcpyDefault);
    if (e == cudaSuccess) e = cudaMalloc((void **)&sc.last, (size_t)nchunks * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&sc.carry, (size_t)nchunks * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&sc.entries, ((size_t)nchunks + 1) * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&sc.base, ((size_t)nchunks + 1) * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&sc.total, 8);
    if (e == cudaSuccess) e = cudaMemset(sc.total, 0, 8);
    if (e != cudaSuccess) return bail(cuda_fail(e, "copyform allocation / copy in", __FILE__, __LINE__));
    SB_LAUNCH(copyform_chunk_kernel<0>, (unsigned)nchunks, kCfThreads, 0, 0, sc.codes, s_rows, s_cols, (int64_t)stride, trans, T,
              cpp, nullptr, sc.last, nullptr, nullptr, nullptr, nullptr, nullptr);
    SB_LAUNCH(copyform_panel_carry_kernel, npanels, kCfThreads, 0, 0, sc.last, cpp, sc.carry);
    SB_LAUNCH(copyform_chunk_kernel<1>, (unsigned)nchunks, kCfThreads, 0, 0, sc.codes, s_rows, s_cols, (int64_t)stride, trans, T,
              cpp, sc.carry, nullptr, sc.entries, sc.total, nullptr, nullptr, nullptr);
    unsigned long long total = 0;
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy(&total, sc.total, 8, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return bail(cuda_fail(e, "copyform count", __FILE__, __LINE__));
    if (total >= (1ull << 31)) return bail(fail(SBLAS_B200_EINVAL, "%llu stream entries do not fit the reference's int32 index bounds", total));
    job->entries = (int64_t)total;
    int rc = exclusive_scan_i32(sc.entries, (int)nchunks, sc.base, 0);
    if (rc != SBLAS_B200_OK) return bail(rc);
    if (total) {
        e = cudaMalloc((void **)&job->d_pos, (size_t)total);
        if (e == cudaSuccess) e = cudaMalloc((void **)&job->d_val, (size_t)total);
        if (e != cudaSuccess) return bail(cuda_fail(e, "copyform stream allocation", __FILE__, __LINE__));
        SB_LAUNCH(copyform_chunk_kernel<2>, (unsigned)nchunks, kCfThreads, 0, 0, sc.codes, s_rows, s_cols, (int64_t)stride, trans,
                  T, cpp, sc.carry, nullptr, nullptr, nullptr, sc.base, job->d_pos, job->d_val);
    }
    // stream offset at every panel start -> the recorded panels (those that emitted anything, sparse-matrix.cc:57)
    std::vector<int32_t> h_base((size_t)nchunks + 1);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy(h_base.data(), sc.base, ((size_t)nchunks + 1) * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return bail(cuda_fail(e, "copyform write", __FILE__, __LINE__));
    for (int p = 0; p < npanels; ++p) {
        const int32_t s = h_base[(size_t)p * cpp], t = h_base[(size_t)(p + 1) * cpp];
        if (t == s) continue;
        job->block_bounds.push_back(0); job->block_bounds.push_back(p * 256);
        job->block_index_bounds.push_back(s); job->block_index_bounds.push_back(t);
    }
    *entries = job->entries;
    *nblocks = (int32_t)(job->block_bounds.size() / 2);
    return SBLAS_B200_OK;
}

int sblas_b200_copyform_finish(sblas_b200_copyform_job *job, uint8_t *pos_index, uint8_t *val_index, int32_t *block_bounds,
                               int32_t *block_index_bounds, const float *val_table, sblas_b200_matrix **mat) {
    if (mat) *mat = nullptr;
    if (!job) return fail(SBLAS_B200_EINVAL, "job is null");
    int rc = SBLAS_B200_OK;
    {
        DeviceScope scope(job->device);
        if (!scope.ok()) rc = SBLAS_B200_ECUDA;
        const size_t nb = job->block_bounds.size();
        if (rc == SBLAS_B200_OK && job->entries && (!pos_index || !val_index)) rc = fail(SBLAS_B200_EINVAL, "null stream output");
        if (rc == SBLAS_B200_OK && nb && (!block_bounds || !block_index_bounds)) rc = fail(SBLAS_B200_EINVAL, "null block output");
        if (rc == SBLAS_B200_OK && job->entries) {
            cudaError_t e = cudaMemcpy(pos_index, job->d_pos, (size_t)job->entries, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess) e = cudaMemcpy(val_index, job->d_val, (size_t)job->entries, cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) rc = cuda_fail(e, "copyform copy out", __FILE__, __LINE__);
        }
        if (rc == SBLAS_B200_OK && nb) {
            std::copy(job->block_bounds.begin(), job->block_bounds.end(), block_bounds);
            std::copy(job->block_index_bounds.begin(), job->block_index_bounds.end(), block_index_bounds);
        }
        if (rc == SBLAS_B200_OK && mat)                  // device layout straight from the device-resident streams
            rc = sblas_b200_upload(job->d_pos, job->d_val, job->entries, val_table, job->T,
                                   nb ? job->block_bounds.data() : nullptr, nb ? job->block_index_bounds.data() : nullptr,
                                   (int32_t)(nb / 2), job->s_rows, job->s_cols, job->device, mat);
    }
    destroy_job(job);
    return rc;
}

void sblas_b200_copyform_abort(sblas_b200_copyform_job *job) { destroy_job(job); }

}  // extern "C"
