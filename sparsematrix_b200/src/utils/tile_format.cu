// tile_format.cu -- second device layout, for the SpMM kernel that stages X through shared memory.
//
// The reference walks its stream panel by panel and, inside a panel, row by row of S, so that
// one transposed A row is reused for every entry of that S row (kernel.cc:780-799, :371-376).
// The CUDA counterpart tiles W = S^T the other way round: a CTA owns a block of W rows and walks
// the X rows in chunks that fit shared memory; inside a chunk every entry of the CTA's rows
// reads its X row from shared memory.  That needs, per (W row, chunk), the entries of that row
// whose column falls in the chunk -- a contiguous piece of the CSR row, since columns ascend.
//
// Layout ("quad rows"): W rows are taken in groups of G = 4 or 8 (one warp), X rows in chunks of
// kTileK = 255.  A warp is 4 slots of 8 lanes; slot s owns row G*g + s (and G*g + 4 + s when
// G = 8) of its group and walks each of them on its own, in column order (the reference's
// accumulation order, kernel.cc:780-796).  One quad row is what the warp consumes in four steps: for every
// slot the next four entries of that slot's current row.  80 bytes:
//     float  w[4][4]   w[s][t] = table[code] of the slot's entry t      (one 128-bit load per slot)
//     uint8  r[4][4]   r[s][t] = column - 255*chunk of that entry, or 255 = "no entry" (skipped)
// For (group, chunk) the quad rows of rows G*g+0..3 come first, then (G = 8) those of rows
// G*g+4..7; each set of four takes as many quad rows as its longest piece needs, shorter
// pieces are filled with "no entry" -- those lanes are predicated off in the kernel and cost
// no shared-memory bandwidth.
// grp_ptr[group*nchunks + chunk] = first quad row of the run;
// grp_nq[group*nchunks + chunk]  = quad rows of rows 0..3 | quad rows of rows 4..7 << 16 (0 when G = 4).
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

// thread per (group, chunk, row-in-group); the 8 threads of one (group, chunk) are neighbours in a warp
__global__ void __launch_bounds__(256)
tile_count_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col, int n, int nchunks, int G,
                  int64_t total, uint32_t *__restrict__ grp_nq, int32_t *__restrict__ grp_blocks) {
    const int64_t idx = blockIdx.x * (int64_t)256 + threadIdx.x;
    const bool in = idx < total;
    const int jj = (int)(idx % G);
    const int64_t gc = idx / G;
    int quads = 0;
    if (in) {
        const int c = (int)(gc % nchunks), g = (int)(gc / nchunks);
        const int j = g * G + jj;
        if (j < n) {
            const int s = row_ptr[j], e = row_ptr[j + 1];
            const int lo = lower_bound_col(col, s, e, c * kTileK);
            const int hi = lower_bound_col(col, lo, e, (c + 1) * kTileK);
            quads = (hi - lo + 3) >> 2;
        }
    }
    // longest piece of each set of four rows: jj = 4*set + slot
    quads = max(quads, __shfl_xor_sync(0xffffffffu, quads, 1));
    quads = max(quads, __shfl_xor_sync(0xffffffffu, quads, 2));
    int other = __shfl_xor_sync(0xffffffffu, quads, 4);
    if (G == 4) other = 0;
    if (in && jj == 0) {
        grp_nq[gc] = (uint32_t)quads | ((uint32_t)other << 16);
        grp_blocks[gc] = quads + other;
    }
}

// thread per (group, chunk, row-in-group): write that row's share of the quad rows
__global__ void __launch_bounds__(256)
tile_fill_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                 const uint8_t *__restrict__ code, const float *__restrict__ table, int n, int nchunks, int G,
                 int64_t total, const uint32_t *__restrict__ grp_nq, const int32_t *__restrict__ grp_ptr,
                 uint8_t *__restrict__ blocks) {
    const int64_t idx = blockIdx.x * (int64_t)256 + threadIdx.x;
    if (idx >= total) return;
    const int jj = (int)(idx % G);
    const int64_t gc = idx / G;
    const uint32_t nq2 = grp_nq[gc];
    const int set = jj >> 2, slot = jj & 3;
    const int nq = set ? (int)(nq2 >> 16) : (int)(nq2 & 0xffffu);
    if (nq == 0) return;
    const int c = (int)(gc % nchunks), g = (int)(gc / nchunks);
    const int j = g * G + jj;
    int lo = 0, len = 0;
    if (j < n) {
        const int s = row_ptr[j], e = row_ptr[j + 1];
        lo = lower_bound_col(col, s, e, c * kTileK);
        len = lower_bound_col(col, lo, e, (c + 1) * kTileK) - lo;
    }
    const int64_t first = (int64_t)grp_ptr[gc] + (set ? (int)(nq2 & 0xffffu) : 0);
    for (int q = 0; q < nq; ++q) {
        uint8_t *blk = blocks + (first + q) * kBlockBytes;
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t r4 = 0xffffffffu;                       // four times "no entry"
        float *wp = &w.x;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int e = 4 * q + t;
            if (e < len) {
                wp[t] = table[code[lo + e]];
                r4 = (r4 & ~(0xffu << (8 * t))) | ((uint32_t)(col[lo + e] - c * kTileK) << (8 * t));
            }
        }
        *reinterpret_cast<float4 *>(blk + slot * 16) = w;
        *reinterpret_cast<uint32_t *>(blk + 64 + slot * 4) = r4;
    }
}

// Build the layout for groups of `rows_per_group` (4 or 8) W rows on `stream`; synchronises it
// (one-time cost, first large-m product of a handle that selects this group size).
int build_tile_format(sblas_b200_matrix *mat, int rows_per_group, cudaStream_t stream) {
    const int G = rows_per_group;
    TileFormat &tf = mat->tile[G == 8];
    if (tf.state != 0) return SBLAS_B200_OK;
    tf.state = -1;
    tf.rows = G;
    if (!mat->csr_sorted || mat->n == 0 || mat->nnz == 0) return SBLAS_B200_OK;
    const int ngroups = (mat->n + G - 1) / G;
    const int nchunks = (mat->k + kTileK - 1) / kTileK;
    const int64_t ngc = (int64_t)ngroups * nchunks;
    // worthwhile only when a (row, chunk) piece holds a few entries on average (otherwise the
    // per-piece bookkeeping outweighs the entries; the gather kernel serves those matrices)
    if (ngc > (1 << 26) || (double)mat->nnz / ((double)mat->n * nchunks) < 1.0) return SBLAS_B200_OK;
    if (const char *env = getenv("SBLAS_B200_NO_TILE")) { if (env[0] == '1') return SBLAS_B200_OK; }
    const int64_t total = ngc * G;
    int32_t *d_grp_blocks = nullptr;
    SB_CUDA(cudaMalloc((void **)&tf.d_grp_nq, (size_t)ngc * 4));
    SB_CUDA(cudaMalloc((void **)&tf.d_grp_ptr, ((size_t)ngc + 1) * 4));
    SB_CUDA(cudaMalloc((void **)&d_grp_blocks, ((size_t)ngc + 1) * 4));
    cudaError_t e = cudaMemsetAsync(d_grp_blocks, 0, ((size_t)ngc + 1) * 4, stream);
    const unsigned grid = (unsigned)((total + 255) / 256);
    if (e == cudaSuccess) {
        SB_LAUNCH(tile_count_kernel, grid, 256, 0, stream, mat->d_row_ptr, mat->d_col, mat->n, nchunks, G, total,
                  tf.d_grp_nq, d_grp_blocks);
        exclusive_scan_i32(d_grp_blocks, (int)ngc, tf.d_grp_ptr, stream);
    }
    int32_t nblocks = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&nblocks, tf.d_grp_ptr + ngc, 4, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    cudaFree(d_grp_blocks);
    if (e != cudaSuccess) return cuda_fail(e, "tile format count", __FILE__, __LINE__);
    // +kTilePad quad rows: the kernel fetches whole ring segments a few segments ahead of use
    const size_t bytes = ((size_t)nblocks + kTilePad) * kBlockBytes;
    SB_CUDA(cudaMalloc((void **)&tf.d_blocks, bytes));
    SB_CUDA(cudaMemsetAsync(tf.d_blocks + (size_t)nblocks * kBlockBytes, 0xff, kTilePad * kBlockBytes, stream));
    SB_LAUNCH(tile_fill_kernel, grid, 256, 0, stream, mat->d_row_ptr, mat->d_col, mat->d_code, mat->d_table, mat->n,
              nchunks, G, total, tf.d_grp_nq, tf.d_grp_ptr, tf.d_blocks);
    SB_CHECK_LAUNCH();
    SB_CUDA(cudaStreamSynchronize(stream));
    tf.ngroups = ngroups;
    tf.nchunks = nchunks;
    tf.nblocks = nblocks;
    mat->format_bytes += (int64_t)bytes + ngc * 4 + (ngc + 1) * 4;
    tf.state = 1;
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200
