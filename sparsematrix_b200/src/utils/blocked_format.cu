// blocked_format.cu -- device-side builder of the column-blocked SpMV layout (DESIGN.md "K-spmv-blocked").
//
// The row-split SpMV kernels gather x[col] straight from L2: with columns spread over an x of many
// megabytes every gather moves its own 32-byte sector and the L2 request rate, not HBM, bounds the
// kernel (power-law case: 0.20 of the HBM roofline).  This layout buys locality with a second copy
// of the entries: sorted by (column block, row, column), two-byte local column and row numbers, so
// that one pass can keep the block's x tile in shared memory and a second pass can add the products
// of a row range out of a plain stream.  The entries of one row inside one column block (a run) sit
// in one quad of slots -- padded if need be -- so that pass 1 can add them and pass 2 sees a row once
// per (block, sub-block) segment; longer runs fill whole quads and are flagged.  Inside a row the order stays column-ascending, which is
// the order the reference adds the terms of one output element in (kernel.cc:780-796).
//
// Three kernels over the CSR of W (nothing is sorted on the host):
//   blocked_count_kernel    warp per row sub-block: entries per column block, number of runs
//   blocked_ptr_kernel      CTA per column block: exclusive scan over the sub-blocks -> slot of every (block, sub-block)
//   blocked_scatter_kernel  warp per row sub-block: stable scatter (stream order kept inside a column block)
// Rows longer than long_threshold stay with the split-row kernels (product.cu) and are left out.
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "device_matrix.h"

namespace sblas_b200 {

namespace {

constexpr int kBlkWarps = 8;            // warps (= row sub-blocks) per CTA of the builder kernels
constexpr int kMaxColBlocks = 256;      // shared-memory counters per warp
constexpr int kSlotAlign = 32;          // a column block starts on a multiple of this many slots (128 B of products)

// Slot rules shared by the counting and the scattering kernel.  A RUN = the entries of one row inside one column block
// (neighbours in the row, columns ascending).  Pass 1 adds the products of a run inside a quad of slots and pass 2 adds
// only the slot that closes it, so a run must not straddle a quad unless it is longer than a quad:
//   run of L <= 4 slots: placed at the cursor, after padding to the next quad if it would not fit the current one;
//   run of L  > 4 slots: starts on a quad, fills whole quads; every quad closes with a partial sum of the row.
__device__ __forceinline__ int run_start(int cursor, int len) {
    if (len > 4 || (cursor & 3) + len > 4) cursor = (cursor + 3) & ~3;
    return cursor;
}

// first index in [lo, hi) whose column is >= bound (columns ascend inside a row)
__device__ __forceinline__ int lower_bound_col(const int32_t *__restrict__ col, int lo, int hi, int64_t bound) {
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if ((int64_t)col[mid] < bound) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(32 * kBlkWarps)
blocked_count_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col, int n, int rs, int nsub, int cbw,
                     int ncb, int long_threshold, int seg_align, int32_t *__restrict__ cnt, int32_t *__restrict__ cb_total,
                     unsigned long long *__restrict__ stats /* [0] runs, [1] entries */) {
    __shared__ int32_t cursor[kBlkWarps][kMaxColBlocks];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sub = blockIdx.x * kBlkWarps + w;
    for (int i = lane; i < kMaxColBlocks; i += 32) cursor[w][i] = 0;
    __syncwarp();
    if (sub >= nsub) return;
    const int r0 = sub * rs, r1 = min(n, r0 + rs);
    unsigned int runs = 0, ents = 0;
    for (int row = r0; row < r1; ++row) {
        const int s = row_ptr[row], e = row_ptr[row + 1];
        if (e - s > long_threshold) continue;
        for (int base = s; base < e; base += 32) {
            const int i = base + lane;
            const bool in = i < e;
            const int cb = in ? col[i] / cbw : -1;
            const bool head = in && (i == s || col[i - 1] / cbw != cb);    // the run's first entry (one run per row and block)
            if (in) ++ents;
            if (head) {
                const int len = lower_bound_col(col, i + 1, e, (int64_t)(cb + 1) * cbw) - i;
                cursor[w][cb] = run_start(cursor[w][cb], len) + len;
                ++runs;
            }
        }
        __syncwarp();                                       // the next row's runs see this row's cursors
    }
    for (int i = lane; i < ncb; i += 32) {
        const int32_t c = (cursor[w][i] + seg_align - 1) & ~(seg_align - 1);   // a sub-segment is a whole number of quads (vector loads), or of 128-byte lines
        cnt[(size_t)i * nsub + sub] = c;
        if (c) atomicAdd(&cb_total[i], c);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        runs += __shfl_xor_sync(0xffffffffu, runs, o);
        ents += __shfl_xor_sync(0xffffffffu, ents, o);
    }
    if (lane == 0) {
        atomicAdd(stats, (unsigned long long)runs);
        atomicAdd(stats + 1, (unsigned long long)ents);
    }
}

// seg_ptr[cb][sub] = cb_start[cb] + slots of (cb, sub' < sub), each sub-segment rounded up to whole quads;
// seg_ptr[cb][nsub] = end of the block's slots
__global__ void __launch_bounds__(256)
blocked_ptr_kernel(const int32_t *__restrict__ cnt, const int32_t *__restrict__ cb_start, int nsub,
                   int32_t *__restrict__ seg_ptr) {
    __shared__ int32_t warp_sum[8];
    __shared__ int32_t chunk_total;
    const int cb = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int32_t *in = cnt + (size_t)cb * nsub;
    int32_t *out = seg_ptr + (size_t)cb * (nsub + 1);
    int32_t carry = cb_start[cb];
    for (int base = 0; base < nsub; base += 256) {
        const int i = base + threadIdx.x;
        const int32_t v = i < nsub ? in[i] : 0;
        int32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_sum[w] = incl;
        __syncthreads();
        if (w == 0) {
            const int32_t t = lane < 8 ? warp_sum[lane] : 0;
            int32_t ti = t;
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {
                const int32_t u = __shfl_up_sync(0xffffffffu, ti, o);
                if (lane >= o) ti += u;
            }
            if (lane < 8) warp_sum[lane] = ti - t;
            if (lane == 7) chunk_total = ti;
        }
        __syncthreads();
        if (i < nsub) out[i] = carry + warp_sum[w] + incl - v;
        carry += chunk_total;
        __syncthreads();
    }
    if (threadIdx.x == 0) out[nsub] = carry;
}

// lcol bit 15: the slot continues the run of the slot before it in the same quad (pass 1 adds the two products).
// lrow: 0xffff = nothing to add from this slot (padding, or a slot inside a run); otherwise the row inside the sub-block
// (bits 0..13) of the run this slot closes, bit 14 set when the run began in an earlier quad (pass 2 then adds it in turn).
__global__ void __launch_bounds__(32 * kBlkWarps)
blocked_scatter_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col, const uint8_t *__restrict__ code,
                       int n, int rs, int nsub, int cbw, int ncb, int long_threshold, const int32_t *__restrict__ seg_ptr,
                       uint16_t *__restrict__ lcol, uint16_t *__restrict__ lrow, uint8_t *__restrict__ ocode) {
    __shared__ int32_t cursor[kBlkWarps][kMaxColBlocks];
    __shared__ int32_t start_of[kBlkWarps][kMaxColBlocks];  // first slot of the current row's run in each block
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sub = blockIdx.x * kBlkWarps + w;
    if (sub >= nsub) return;
    for (int i = lane; i < ncb; i += 32) cursor[w][i] = seg_ptr[(size_t)i * (nsub + 1) + sub];
    __syncwarp();
    const int r0 = sub * rs, r1 = min(n, r0 + rs);
    for (int row = r0; row < r1; ++row) {
        const int s = row_ptr[row], e = row_ptr[row + 1];
        if (e - s > long_threshold) continue;
        for (int base = s; base < e; base += 32) {
            const int i = base + lane;
            const bool in = i < e;
            const int c = in ? col[i] : 0;
            const int cb = in ? c / cbw : 0;
            const bool head = in && (i == s || col[i - 1] / cbw != cb);
            if (head) {
                const int len = lower_bound_col(col, i + 1, e, (int64_t)(cb + 1) * cbw) - i;
                const int st = run_start(cursor[w][cb], len);   // slots are relative to nothing: seg_ptr is quad aligned
                start_of[w][cb] = st;
                cursor[w][cb] = st + len;
            }
            __syncwarp();
            if (in) {
                const int first = head ? i : lower_bound_col(col, s, i, (int64_t)cb * cbw);   // index of the run's first entry
                const int st = start_of[w][cb];
                const int d = st + (i - first);
                const bool last = i + 1 == e || col[i + 1] / cbw != cb;
                uint16_t lc = (uint16_t)(c - cb * cbw);
                if (i > first && (d & 3) != 0) lc |= 0x8000u;
                uint16_t lr = 0xffffu;
                if (last || (d & 3) == 3) lr = (uint16_t)((row - r0) | (((d >> 2) != (st >> 2)) ? 0x4000 : 0));
                lcol[d] = lc;
                lrow[d] = lr;
                ocode[d] = code[i];
            }
            __syncwarp();
        }
    }
}

int env_int(const char *name, int fallback) {
    const char *e = getenv(name);
    return (e && *e) ? atoi(e) : fallback;
}

}  // namespace

// When the layout pays: x far beyond the L1s (>= 2^21 columns = 8 MB) and enough entries to fill
// the machine with row sub-blocks of useful size (>= 2^26: every warp of pass 2 walks all column blocks one after
// the other, about a microsecond each, whatever its share of the rows -- a half-size slab of the 2^23-row power-law
// matrix takes 332 us this way and 305 us with the length-sorted row kernel).  SBLAS_B200_SPMV_BLOCKED = 0 | 1 forces
// it off / on (tests build it on small matrices; _CBW and _RS then shrink the geometry).
int build_blocked_format(sblas_b200_matrix *mat) {
    BlockedFormat &bf = mat->blocked;
    if (bf.state != 0) return SBLAS_B200_OK;
    bf.state = -1;
    const int mode = env_int("SBLAS_B200_SPMV_BLOCKED", -1);
    if (mode == 0 || !mat->csr_sorted || mat->n < 1 || mat->k < 1 || mat->nnz < 1) return SBLAS_B200_OK;
    if (mode != 1 && (mat->k < (1 << 21) || mat->nnz < ((int64_t)1 << 26))) return SBLAS_B200_OK;
    int cbw = env_int("SBLAS_B200_BLOCKED_CBW", 32768);
    cbw = std::min(32768, std::max(4, cbw & ~3));           // 15 bits of local column, bit 15 of lcol is a flag
    const int ncb = (int)(((int64_t)mat->k + cbw - 1) / cbw);
    if (ncb > kMaxColBlocks) return SBLAS_B200_OK;
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mat->device) != cudaSuccess || sms <= 0) sms = 148;
    // pass 2 runs one warp per sub-block, 16 per CTA, two CTAs per SM: every sub-block resident at once when the matrix
    // is large enough for sub-segments of >= 96 entries (a warp adds 128 slots per step); a CTA's row sums must fit
    // half an SM's shared memory (16 * rs floats <= 112 KB)
    int64_t want_sub = std::min<int64_t>((int64_t)2 * sms * 16, mat->nnz / ((int64_t)ncb * 96));
    want_sub = std::max<int64_t>(16, want_sub);
    int rs = (int)(((int64_t)mat->n + want_sub - 1) / want_sub);
    rs = std::min(rs, kBlockedMaxRows);
    if (const int v = env_int("SBLAS_B200_BLOCKED_RS", 0)) rs = std::min(kBlockedMaxRows, std::max(1, v));
    rs = std::max(rs, 1);
    const int nsub = (int)(((int64_t)mat->n + rs - 1) / rs);

    struct Scratch {
        int32_t *cnt = nullptr, *cb_total = nullptr;
        unsigned long long *stats = nullptr;
        ~Scratch() { cudaFree(cnt); cudaFree(cb_total); cudaFree(stats); }
    } sc;
    SB_CUDA(cudaMalloc((void **)&sc.cnt, (size_t)ncb * nsub * 4));
    SB_CUDA(cudaMalloc((void **)&sc.cb_total, (size_t)ncb * 4));
    SB_CUDA(cudaMalloc((void **)&sc.stats, 16));
    SB_CUDA(cudaMemset(sc.cb_total, 0, (size_t)ncb * 4));
    SB_CUDA(cudaMemset(sc.stats, 0, 16));
    const int grid = (nsub + kBlkWarps - 1) / kBlkWarps;
    int seg_align = env_int("SBLAS_B200_BLOCKED_ALIGN", 4);
    if (seg_align != 4 && seg_align != 8 && seg_align != 16 && seg_align != 32) seg_align = 4;
    SB_LAUNCH(blocked_count_kernel, grid, 32 * kBlkWarps, 0, 0, mat->d_row_ptr, mat->d_col, mat->n, rs, nsub, cbw, ncb,
              mat->long_threshold, seg_align, sc.cnt, sc.cb_total, sc.stats);
    SB_CHECK_LAUNCH();
    std::vector<int32_t> cb_total(ncb), cb_start(ncb + 1, 0);
    unsigned long long stats[2] = {0, 0};
    SB_CUDA(cudaMemcpy(cb_total.data(), sc.cb_total, (size_t)ncb * 4, cudaMemcpyDeviceToHost));
    SB_CUDA(cudaMemcpy(stats, sc.stats, 16, cudaMemcpyDeviceToHost));
    bf.runs = (int64_t)stats[0];
    bf.entries = (int64_t)stats[1];
    if (bf.entries == 0) return SBLAS_B200_OK;              // only long rows: nothing for this layout
    // Clustered columns (bands, blocks): a row's entries share a column block, the row kernels' gathers hit L1
    // and 13 bytes per entry would be a loss.  Random columns give about one run per entry.
    if (mode != 1 && (double)bf.runs < 0.5 * (double)bf.entries) return SBLAS_B200_OK;
    int64_t total = 0;
    for (int cb = 0; cb < ncb; ++cb) {
        cb_start[cb] = (int32_t)total;
        total += ((int64_t)cb_total[cb] + kSlotAlign - 1) / kSlotAlign * kSlotAlign;
        if (total >= ((int64_t)1 << 31) - 4096) return SBLAS_B200_OK;    // slots are int32
    }
    cb_start[ncb] = (int32_t)total;

    auto give_up = [&](cudaError_t e) {                     // no room for the second copy: the row kernels stay
        cudaFree(bf.d_lcol); cudaFree(bf.d_lrow); cudaFree(bf.d_code); cudaFree(bf.d_seg_ptr); cudaFree(bf.d_cb_start);
        cudaFree(bf.d_prod);
        bf = BlockedFormat();
        bf.state = -1;
        if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); return (int)SBLAS_B200_OK; }
        return cuda_fail(e, "blocked SpMV layout", __FILE__, __LINE__);
    };
    cudaError_t e = cudaMalloc((void **)&bf.d_lcol, (size_t)total * 2);
    if (e == cudaSuccess) e = cudaMalloc((void **)&bf.d_lrow, (size_t)total * 2);
    if (e == cudaSuccess) e = cudaMalloc((void **)&bf.d_code, (size_t)total);
    if (e == cudaSuccess) e = cudaMalloc((void **)&bf.d_prod, (size_t)total * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&bf.d_seg_ptr, (size_t)ncb * (nsub + 1) * 4);
    if (e == cudaSuccess) e = cudaMalloc((void **)&bf.d_cb_start, ((size_t)ncb + 1) * 4);
    if (e == cudaSuccess) e = cudaMemset(bf.d_lcol, 0, (size_t)total * 2);
    if (e == cudaSuccess) e = cudaMemset(bf.d_lrow, 0xff, (size_t)total * 2);  // padding slots: row 0xffff = no row
    if (e == cudaSuccess) e = cudaMemset(bf.d_code, 0xff, (size_t)total);      // padding slots: no codebook entry
    if (e == cudaSuccess) e = cudaMemcpy(bf.d_cb_start, cb_start.data(), ((size_t)ncb + 1) * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return give_up(e);
    SB_LAUNCH(blocked_ptr_kernel, ncb, 256, 0, 0, sc.cnt, bf.d_cb_start, nsub, bf.d_seg_ptr);
    SB_LAUNCH(blocked_scatter_kernel, grid, 32 * kBlkWarps, 0, 0, mat->d_row_ptr, mat->d_col, mat->d_code, mat->n, rs, nsub,
              cbw, ncb, mat->long_threshold, bf.d_seg_ptr, bf.d_lcol, bf.d_lrow, bf.d_code);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return give_up(e);
    bf.cbw = cbw; bf.ncb = ncb; bf.rs = rs; bf.nsub = nsub; bf.total = total;
    bf.state = 1;
    mat->format_bytes += total * 9 + (int64_t)ncb * (nsub + 1) * 4 + (ncb + 1) * 4;
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200
