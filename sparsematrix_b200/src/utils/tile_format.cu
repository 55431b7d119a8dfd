// tile_format.cu -- second device layout, for the SpMM kernel that stages X through shared memory.
//
// The reference walks its stream panel by panel and, inside a panel, row by row of S, so that
// one transposed A row is reused for every entry of that S row (kernel.cc:780-799, :371-376).
// The CUDA counterpart tiles W = S^T the other way round: a CTA owns a block of W rows and walks
// the X rows in chunks that fit shared memory; inside a chunk every entry of the CTA's rows
// reads its X row from shared memory.  That needs, per (W row, chunk), the entries of that row
// whose column falls in the chunk -- a contiguous piece of the CSR row, since columns ascend.
//
// Layout ("blocks"): W rows are taken in groups of kTileRows = 8, X rows in chunks of
// kTileK = 255.  For each (group, chunk), the pieces of the 8 rows follow one another, each
// padded to whole blocks of 16 entries.  A block is 80 bytes:
//     float  w[4][4]   w[s][t] = table[code] of entry 4*t + s      (s = warp slot, t = step)
//     uint8  r[4][4]   r[s][t] = column - 255*chunk of that entry  (255 = the zero row, padding)
// so that one 128-bit load gives a slot its four next weights and one 32-bit load its four
// next X-row offsets.  Padding entries are (w = 0, r = 255): they multiply an all-zero row of
// shared memory and so never touch real X data.
// grp_ptr[group*nchunks + chunk] = first block of the run; seg_len[(group*nchunks + chunk)*8 + row]
// = number of real entries of that row in the chunk.
#include <stdlib.h>

#include <algorithm>

#include "device_matrix.h"

namespace sblas_b200 {

__device__ __forceinline__ int lower_bound_col(const int32_t *__restrict__ col, int lo, int hi, int key) {
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (col[mid] < key) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

// thread per (group, chunk, row-in-group): entries of that row inside the chunk; blocks per (group, chunk)
__global__ void __launch_bounds__(256)
tile_count_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col, int n, int nchunks,
                  int64_t total, uint16_t *__restrict__ seg_len, int32_t *__restrict__ grp_blocks) {
    const int64_t idx = blockIdx.x * (int64_t)256 + threadIdx.x;
    if (idx >= total) return;
    const int jj = (int)(idx % kTileRows);
    const int64_t gc = idx / kTileRows;
    const int c = (int)(gc % nchunks), g = (int)(gc / nchunks);
    const int j = g * kTileRows + jj;
    int len = 0;
    if (j < n) {
        const int s = row_ptr[j], e = row_ptr[j + 1];
        const int lo = lower_bound_col(col, s, e, c * kTileK);
        const int hi = lower_bound_col(col, lo, e, (c + 1) * kTileK);
        len = hi - lo;
    }
    seg_len[idx] = (uint16_t)len;
    if (len) atomicAdd(&grp_blocks[gc], (len + kBlockEntries - 1) / kBlockEntries);
}

// thread per (group, chunk, row-in-group): write that row's blocks
__global__ void __launch_bounds__(256)
tile_fill_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                 const uint8_t *__restrict__ code, const float *__restrict__ table, int n, int nchunks,
                 int64_t total, const uint16_t *__restrict__ seg_len, const int32_t *__restrict__ grp_ptr,
                 uint8_t *__restrict__ blocks) {
    const int64_t idx = blockIdx.x * (int64_t)256 + threadIdx.x;
    if (idx >= total) return;
    const int jj = (int)(idx % kTileRows);
    const int64_t gc = idx / kTileRows;
    const int len = seg_len[idx];
    if (len == 0) return;
    const int c = (int)(gc % nchunks), g = (int)(gc / nchunks);
    const int j = g * kTileRows + jj;
    int64_t blk = grp_ptr[gc];
    for (int q = 0; q < jj; ++q) blk += (seg_len[gc * kTileRows + q] + kBlockEntries - 1) / kBlockEntries;
    const int lo = lower_bound_col(col, row_ptr[j], row_ptr[j + 1], c * kTileK);
    const int nblk = (len + kBlockEntries - 1) / kBlockEntries;
    for (int b = 0; b < nblk; ++b) {
        float *w = reinterpret_cast<float *>(blocks + (blk + b) * kBlockBytes);
        uint8_t *r = blocks + (blk + b) * kBlockBytes + 64;
        for (int q = 0; q < kBlockEntries; ++q) {
            const int e = b * kBlockEntries + q;          // entry 4*t + s  ->  w[s][t]
            const int s = q & 3, t = q >> 2;
            float wv = 0.0f;
            int rv = kTileK;                              // the zero row
            if (e < len) {
                wv = table[code[lo + e]];
                rv = col[lo + e] - c * kTileK;
            }
            w[s * 4 + t] = wv;
            r[s * 4 + t] = (uint8_t)rv;
        }
    }
}

// Build on `stream`; synchronises it (one-time cost, first large-m product of a handle).
int build_tile_format(sblas_b200_matrix *mat, cudaStream_t stream) {
    if (mat->tile_state != 0) return SBLAS_B200_OK;
    mat->tile_state = -1;
    if (!mat->csr_sorted || mat->n == 0 || mat->nnz == 0) return SBLAS_B200_OK;
    const int ngroups = (mat->n + kTileRows - 1) / kTileRows;
    const int nchunks = (mat->k + kTileK - 1) / kTileK;
    const int64_t ngc = (int64_t)ngroups * nchunks;
    // worthwhile only when a (row, chunk) piece holds a few entries on average (otherwise the
    // per-piece bookkeeping outweighs the entries; the gather kernel serves those matrices)
    if (ngc > (1 << 26) || (double)mat->nnz / ((double)ngc * kTileRows) < 1.0) return SBLAS_B200_OK;
    if (const char *env = getenv("SBLAS_B200_NO_TILE")) { if (env[0] == '1') return SBLAS_B200_OK; }
    const int64_t total = ngc * kTileRows;
    int32_t *d_grp_blocks = nullptr;
    SB_CUDA(cudaMalloc((void **)&mat->d_seg_len, (size_t)total * 2));
    SB_CUDA(cudaMalloc((void **)&mat->d_grp_ptr, ((size_t)ngc + 1) * 4));
    SB_CUDA(cudaMalloc((void **)&d_grp_blocks, ((size_t)ngc + 1) * 4));
    cudaError_t e = cudaMemsetAsync(d_grp_blocks, 0, ((size_t)ngc + 1) * 4, stream);
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (e == cudaSuccess) {
        SB_LAUNCH(tile_count_kernel, grid, 256, 0, stream, mat->d_row_ptr, mat->d_col, mat->n, nchunks, total,
                  mat->d_seg_len, d_grp_blocks);
        exclusive_scan_i32(d_grp_blocks, (int)ngc, mat->d_grp_ptr, stream);
    }
    int32_t nblocks = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&nblocks, mat->d_grp_ptr + ngc, 4, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    cudaFree(d_grp_blocks);
    if (e != cudaSuccess) return cuda_fail(e, "tile format count", __FILE__, __LINE__);
    // +2 blocks: the kernel's look-ahead loads run two blocks past the last real one
    const size_t bytes = ((size_t)nblocks + 2) * kBlockBytes;
    SB_CUDA(cudaMalloc((void **)&mat->d_tile_blocks, bytes));
    SB_CUDA(cudaMemsetAsync(mat->d_tile_blocks + (size_t)nblocks * kBlockBytes, 0, 2 * kBlockBytes, stream));
    SB_LAUNCH(tile_fill_kernel, grid, 256, 0, stream, mat->d_row_ptr, mat->d_col, mat->d_code, mat->d_table, mat->n,
              nchunks, total, mat->d_seg_len, mat->d_grp_ptr, mat->d_tile_blocks);
    SB_CHECK_LAUNCH();
    SB_CUDA(cudaStreamSynchronize(stream));
    mat->tile_ngroups = ngroups;
    mat->tile_nchunks = nchunks;
    mat->tile_nblocks = nblocks;
    mat->format_bytes += (int64_t)bytes + total * 2 + (ngc + 1) * 4;
    mat->tile_state = 1;
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200
