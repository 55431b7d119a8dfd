// device_format.cu -- format upload and device allocation (north_star: "Format
// upload and device allocation move into src/utils").
//
// Input: the compressed members of sblas::SparseMatrix<uint8,uint8,float>
// (reference src/sparse/sparse-matrix.h:45-52): per 256-column panel a byte stream
// of position deltas and a byte stream of codebook indices, fillers bridging gaps
// > 255 (sparse-matrix.cc:46-51).  Output: CSR of W = S^T with 1-byte codes, the
// layout the product kernels read (DESIGN.md "Data layout in HBM").
//
// The conversion runs on the device and restates the decode loop of
// kernel.cc:780-796 as data-parallel steps:
//   K-up1  seg_delta_sum      per 2048-entry segment: sum of the deltas
//   K-up2  panel_seg_scan     per panel: exclusive scan of the segment sums -> segment start position
//   K-up3  seg_count          per segment: warp scan of deltas -> (r, jj); drop fillers; 256-bin histogram
//   K-up4  panel_bin_scan     per panel, per bin: exclusive scan over segments -> stable offsets, row lengths
//   K-up5  row_ptr_scan       exclusive scan of row lengths -> row_ptr
//   K-up6  seg_scatter        per segment: stable scatter (match_any ranks) -> col[], code[]
// Inside one W row the entries keep stream order = ascending r, which is the
// reference's accumulation order for that output element.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "device_matrix.h"

namespace sblas_b200 {

static thread_local char t_error[512] = "";
std::atomic<int64_t> g_kernel_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_error, sizeof(t_error), fmt, ap);
    va_end(ap);
}

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_error, sizeof(t_error), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
    const char *base = strrchr(file, '/');
    snprintf(t_error, sizeof(t_error), "CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e),
             base ? base + 1 : file, line, what);
    return e == cudaErrorMemoryAllocation ? SBLAS_B200_ENOMEM : SBLAS_B200_ECUDA;
}

DeviceScope::DeviceScope(int device) : prev_(-1), ok_(false) {
    if (cudaGetDevice(&prev_) != cudaSuccess) { prev_ = -1; }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) { cuda_fail(e, "cudaSetDevice", __FILE__, __LINE__); return; }
    ok_ = true;
}
DeviceScope::~DeviceScope() {
    if (prev_ >= 0) cudaSetDevice(prev_);
}

void note_product(sblas_b200_matrix *mat, const char *kernel, int m, double bytes_per_nnz, int64_t index_bytes) {
    snprintf(mat->last_kernel, sizeof(mat->last_kernel), "%s", kernel);
    mat->last_m = m;
    mat->last_bytes_per_nnz = bytes_per_nnz;
    mat->last_bytes = (int64_t)(bytes_per_nnz * (double)mat->nnz + 0.5) + index_bytes + (int64_t)4 * mat->k * m + (int64_t)8 * mat->n * m;
}

int ensure_buffer(float **buf, size_t *have, size_t need) {
    if (need <= *have) return SBLAS_B200_OK;
    if (*buf) { cudaFree(*buf); *buf = nullptr; *have = 0; }
    size_t want = need + need / 4;          // grow with slack: repeated calls with growing m do not thrash
    SB_CUDA(cudaMalloc((void **)buf, want));
    *have = want;
    return SBLAS_B200_OK;
}

// ---------------------------------------------------------------------------------------------
// upload kernels
// ---------------------------------------------------------------------------------------------
constexpr int kSegEntries = 2048;   // stream entries per segment (one warp each)
constexpr int kWarpsPerCta = 8;

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
    }
    return v;
}

// K-up1
template <typename PosT>
__global__ void __launch_bounds__(32 * kWarpsPerCta)
seg_delta_sum_kernel(const PosT *__restrict__ pos, const int32_t *__restrict__ seg_begin,
                     const int32_t *__restrict__ seg_end, int nseg, uint32_t *__restrict__ seg_pos0) {
    const int seg = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (seg >= nseg) return;
    uint32_t s = 0;
    for (int e = seg_begin[seg] + lane; e < seg_end[seg]; e += 32) s += pos[e];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) seg_pos0[seg] = s;
}

// K-up2: one thread per panel; the number of segments per panel is small (entries/2048).
__global__ void panel_seg_scan_kernel(const int32_t *__restrict__ panel_seg_first, int npanels,
                                      uint32_t *__restrict__ seg_pos0) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npanels) return;
    uint32_t run = 0;                                   // prev_index = 0 at panel start (sparse-matrix.cc:36)
    for (int s = panel_seg_first[p]; s < panel_seg_first[p + 1]; ++s) {
        const uint32_t d = seg_pos0[s];
        seg_pos0[s] = run;
        run += d;
    }
}

// K-up3
template <typename PosT>
__global__ void __launch_bounds__(32 * kWarpsPerCta)
seg_count_kernel(const PosT *__restrict__ pos, const uint8_t *__restrict__ code, int T,
                 const int32_t *__restrict__ seg_begin, const int32_t *__restrict__ seg_end,
                 const int32_t *__restrict__ seg_panel, const int32_t *__restrict__ panel_col_off,
                 const uint32_t *__restrict__ seg_pos0, int nseg, int n, int k, int32_t *__restrict__ seg_hist,
                 unsigned long long *__restrict__ fillers /* [0] fillers, [1] entries outside the matrix */) {
    __shared__ int32_t hist[kWarpsPerCta][kPanelWidth];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seg = blockIdx.x * kWarpsPerCta + w;
    for (int i = lane; i < kPanelWidth; i += 32) hist[w][i] = 0;
    __syncwarp();
    if (seg >= nseg) return;
    uint32_t running = seg_pos0[seg];
    int my_fill = 0, my_bad = 0;
    const int col_off = panel_col_off[seg_panel[seg]];
    const int sb = seg_begin[seg], se = seg_end[seg];
    for (int base = sb; base < se; base += 32) {
        const int e = base + lane;
        const bool in = e < se;
        const uint32_t d = in ? pos[e] : 0u;
        const int cd = in ? code[e] : 0;
        const uint32_t incl = warp_inclusive_scan(d, lane);
        const uint32_t p = running + incl;              // pos_offset += ppos[i]   (kernel.cc:781)
        running += __shfl_sync(0xffffffffu, incl, 31);
        if (in) {
            if (cd >= T) ++my_fill;                     // filler: code >= valid_table_size (kernel.cc:782)
            else if (col_off + (int)(p & kPanelMask) >= n || (p >> kPanelShift) >= (uint32_t)k) ++my_bad;
            else atomicAdd(&hist[w][p & kPanelMask], 1);
        }
    }
    __syncwarp();
    for (int i = lane; i < kPanelWidth; i += 32) seg_hist[(size_t)seg * kPanelWidth + i] = hist[w][i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_fill += __shfl_xor_sync(0xffffffffu, my_fill, o);
    if (lane == 0 && my_fill) atomicAdd(fillers, (unsigned long long)my_fill);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_bad += __shfl_xor_sync(0xffffffffu, my_bad, o);
    if (lane == 0 && my_bad) atomicAdd(fillers + 1, (unsigned long long)my_bad);
}

// K-up4: CTA per panel, thread per bin (= column inside the panel = local W row)
__global__ void __launch_bounds__(kPanelWidth)
panel_bin_scan_kernel(const int32_t *__restrict__ panel_seg_first, const int32_t *__restrict__ panel_col_off,
                      int32_t *__restrict__ seg_hist, int32_t *__restrict__ row_len, int n) {
    const int p = blockIdx.x, jj = threadIdx.x;
    int32_t run = 0;
    for (int s = panel_seg_first[p]; s < panel_seg_first[p + 1]; ++s) {
        const int32_t c = seg_hist[(size_t)s * kPanelWidth + jj];
        seg_hist[(size_t)s * kPanelWidth + jj] = run;
        run += c;
    }
    const int j = panel_col_off[p] + jj;
    if (j < n) row_len[j] = run;
}

// K-up5: single CTA, chunked exclusive scan; row_ptr[n] = total
__global__ void __launch_bounds__(1024)
row_ptr_scan_kernel(const int32_t *__restrict__ row_len, int n, int32_t *__restrict__ row_ptr) {
    __shared__ uint32_t warp_off[32];
    __shared__ uint32_t chunk_total;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t carry = 0;                                  // every thread tracks the same running total
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < n ? (uint32_t)row_len[i] : 0u;
        const uint32_t incl = warp_inclusive_scan(v, lane);
        if (lane == 31) warp_off[w] = incl;
        __syncthreads();
        if (w == 0) {
            const uint32_t t = warp_off[lane];
            const uint32_t ti = warp_inclusive_scan(t, lane);
            warp_off[lane] = ti - t;
            if (lane == 31) chunk_total = ti;
        }
        __syncthreads();
        if (i < n) row_ptr[i] = (int32_t)(carry + warp_off[w] + incl - v);
        carry += chunk_total;
        __syncthreads();
    }
    if (threadIdx.x == 0) row_ptr[n] = (int32_t)carry;
}

// K-up6
template <typename PosT>
__global__ void __launch_bounds__(32 * kWarpsPerCta)
seg_scatter_kernel(const PosT *__restrict__ pos, const uint8_t *__restrict__ code, int T,
                   const int32_t *__restrict__ seg_begin, const int32_t *__restrict__ seg_end,
                   const int32_t *__restrict__ seg_panel, const uint32_t *__restrict__ seg_pos0, int nseg,
                   const int32_t *__restrict__ panel_col_off, const int32_t *__restrict__ seg_hist,
                   const int32_t *__restrict__ row_ptr, int n, int k, int32_t *__restrict__ out_col,
                   uint8_t *__restrict__ out_code) {
    __shared__ int32_t cursor[kWarpsPerCta][kPanelWidth];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seg = blockIdx.x * kWarpsPerCta + w;
    if (seg >= nseg) return;
    const int col_off = panel_col_off[seg_panel[seg]];
    for (int i = lane; i < kPanelWidth; i += 32) cursor[w][i] = seg_hist[(size_t)seg * kPanelWidth + i];
    __syncwarp();
    uint32_t running = seg_pos0[seg];
    const int sb = seg_begin[seg], se = seg_end[seg];
    for (int base = sb; base < se; base += 32) {
        const int e = base + lane;
        const bool in = e < se;
        const uint32_t d = in ? pos[e] : 0u;
        const int cd = in ? code[e] : 0;
        const uint32_t incl = warp_inclusive_scan(d, lane);
        const uint32_t p = running + incl;
        running += __shfl_sync(0xffffffffu, incl, 31);
        const int jj = (int)(p & kPanelMask);
        const bool real = in && cd < T && col_off + jj < n && (p >> kPanelShift) < (uint32_t)k;
        // lanes holding the same W row, in lane (= stream) order -> stable rank
        const unsigned peers = __match_any_sync(0xffffffffu, real ? jj : (kPanelWidth + lane));
        const int rank = __popc(peers & ((1u << lane) - 1u));
        int off = 0;
        if (real) off = cursor[w][jj];
        __syncwarp();
        if (real && lane == 31 - __clz(peers)) cursor[w][jj] = off + __popc(peers);
        __syncwarp();
        if (real) {
            const int dest = row_ptr[col_off + jj] + off + rank;
            out_col[dest] = (int32_t)(p >> kPanelShift);   // row = pos_offset >> block_col_shift (kernel.cc:783)
            out_code[dest] = (uint8_t)cd;
        }
    }
}

int exclusive_scan_i32(const int32_t *d_in, int n, int32_t *d_out, cudaStream_t stream) {
    SB_LAUNCH(row_ptr_scan_kernel, 1, 1024, 0, stream, d_in, n, d_out);
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

// CSR validation for direct uploads: every column index inside [0, k), every code < T.
__global__ void csr_validate_kernel(const int32_t *__restrict__ col, const uint8_t *__restrict__ code, int64_t nnz,
                                    int k, int T, int *__restrict__ bad) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x)
        if ((unsigned)col[i] >= (unsigned)k || code[i] >= T) atomicOr(bad, 1);
}

// bit 1 of *bad: some row's columns are not strictly ascending (the tile format then stays off)
__global__ void csr_sorted_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col, int n,
                                  int *__restrict__ bad) {
    const int j = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (j >= n) return;
    for (int e = row_ptr[j] + 1 + lane; e < row_ptr[j + 1]; e += 32)
        if (col[e] <= col[e - 1]) atomicOr(bad, 2);
}

// Sorted SpMV layout: warp per sorted position copies its row, padding the last block (col 0, code 255).
__global__ void __launch_bounds__(256)
sorted_copy_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col, const uint8_t *__restrict__ code,
                   const int32_t *__restrict__ order, const int32_t *__restrict__ blk_ptr, int n_ordered,
                   int32_t *__restrict__ scol, uint8_t *__restrict__ scode) {
    const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n_ordered) return;
    const int row = order[i];
    const int s = row_ptr[row], len = row_ptr[row + 1] - s;
    const int64_t d = (int64_t)blk_ptr[i] * 4;
    const int padded = (len + 3) & ~3;
    for (int e = lane; e < padded; e += 32) {
        scol[d + e] = e < len ? col[s + e] : 0;
        scode[d + e] = e < len ? code[s + e] : (uint8_t)255;
    }
}

// K-decode: CSR(W) -> dense, the function of CopyTo (sparse-matrix.cc:101-137).  Warp per W row.
__global__ void __launch_bounds__(256)
decode_dense_kernel(const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                    const uint8_t *__restrict__ code, const float *__restrict__ table, int n,
                    float *__restrict__ dense, int stride, int trans) {
    const int j = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (j >= n) return;
    for (int e = row_ptr[j] + lane; e < row_ptr[j + 1]; e += 32) {
        const int r = col[e];
        const float v = table[code[e]];
        if (trans) dense[(size_t)j * stride + r] = v;
        else dense[(size_t)r * stride + j] = v;
    }
}

// ---------------------------------------------------------------------------------------------
// host side of the upload
// ---------------------------------------------------------------------------------------------
static int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

int plan_spmv(sblas_b200_matrix *mat, const int32_t *h_row_ptr) {
    const int n = mat->n;
    int64_t nonempty = 0;
    int32_t max_len = 0;
    for (int b = 0; b < 33; ++b) mat->row_len_hist[b] = 0;
    for (int j = 0; j < n; ++j) {
        const int32_t len = h_row_ptr[j + 1] - h_row_ptr[j];
        if (len > max_len) max_len = len;
        int b = 0;
        if (len > 0) { ++nonempty; b = 32 - __builtin_clz((unsigned)len); }
        ++mat->row_len_hist[b];
    }
    mat->max_row_nnz = max_len;
    const double mean = nonempty ? (double)mat->nnz / (double)nonempty : 0.0;
    // each lane consumes one aligned block of 4 non-zeros per iteration: aim at ~2 iterations for a typical row
    // measured on config 4 (mean 16): 2 lanes beat 4 and 8 -- the row kernel is bound by the L1TEX
    // rate of the scattered x gathers, and fewer lanes per row mean fewer idle lanes on short rows
    int lanes = next_pow2((int)((mean + 7.0) / 8.0));
    lanes = std::min(32, std::max(2, lanes));
    if (const char *env = getenv("SBLAS_B200_SPMV_LANES")) {
        const int v = atoi(env);
        if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16 || v == 32) lanes = v;
    }
    mat->spmv_lanes = lanes;
    // a sub-warp streams rows up to 512 entries per lane; at 256 per lane the 16384^2 / 50 % case had half
    // of its rows (about 8192 entries) on the split-row path and ran 207 instead of 158 us
    // -- with few lanes per row (short mean, i.e. skewed lengths) a row of 1024 entries is a chain of 128 dependent
    // iterations of one lane pair, ~130 us whatever else the kernel does: a 1/8 slab of the power-law matrix ran 194 us
    // with the threshold at 1024 and 108 us at 256 (r02); such rows go to the split-row path earlier
    mat->long_threshold = lanes >= 8 ? 512 * lanes : std::max(256, 128 * lanes);
    if (const char *env = getenv("SBLAS_B200_LONG_THRESHOLD")) {
        const int v = atoi(env);
        if (v > 0) mat->long_threshold = v;
    }
    std::vector<int32_t> long_rows, seg_ptr, seg_b, seg_e;
    seg_ptr.push_back(0);
    for (int j = 0; j < n; ++j) {
        const int32_t s = h_row_ptr[j], e = h_row_ptr[j + 1];
        if (e - s <= mat->long_threshold) continue;
        long_rows.push_back(j);
        for (int32_t b = s; b < e; b += kLongSegment) {
            seg_b.push_back(b);
            seg_e.push_back(std::min(e, b + kLongSegment));
        }
        seg_ptr.push_back((int32_t)seg_b.size());
    }
    mat->n_long = (int32_t)long_rows.size();
    mat->n_long_seg = (int32_t)seg_b.size();
    // Skewed row lengths: a warp is as slow as its longest row, so rows are visited in order of
    // decreasing length (counting sort) and a warp's rows are about equally long.
    bool sort_rows = n >= 65536 && (double)max_len >= 4.0 * std::max(1.0, mean);
    if (const char *env = getenv("SBLAS_B200_SORT_ROWS")) sort_rows = env[0] == '1';
    // x of many megabytes and scattered columns: the column-blocked copy replaces the sorted one (blocked_format.cu)
    {
        const int brc = build_blocked_format(mat);
        if (brc != SBLAS_B200_OK) return brc;
        if (mat->blocked.state == 1) sort_rows = false;
    }
    if (sort_rows) {
        const int32_t cap = mat->long_threshold;
        std::vector<int32_t> count((size_t)cap + 2, 0), order;
        for (int j = 0; j < n; ++j) {
            const int32_t len = h_row_ptr[j + 1] - h_row_ptr[j];
            if (len <= cap) ++count[cap - len + 1];              // bucket 0 = longest
        }
        for (int32_t b = 0; b <= cap; ++b) count[b + 1] += count[b];
        order.resize((size_t)count[cap + 1]);
        for (int j = 0; j < n; ++j) {
            const int32_t len = h_row_ptr[j + 1] - h_row_ptr[j];
            if (len <= cap) order[(size_t)count[cap - len]++] = j;
        }
        mat->n_ordered = (int32_t)order.size();
        if (mat->n_ordered) {
            std::vector<int32_t> blk_ptr(order.size() + 1, 0), slen(order.size());
            int64_t blocks = 0;
            for (size_t i = 0; i < order.size(); ++i) {
                blk_ptr[i] = (int32_t)blocks;
                slen[i] = h_row_ptr[order[i] + 1] - h_row_ptr[order[i]];
                blocks += (slen[i] + 3) >> 2;
            }
            blk_ptr[order.size()] = (int32_t)blocks;
            if (blocks >= ((int64_t)1 << 31)) return fail(SBLAS_B200_EINVAL, "sorted SpMV layout does not fit int32 blocks");
            mat->sorted_blocks = blocks;
            SB_CUDA(cudaMalloc((void **)&mat->d_row_order, order.size() * 4));
            SB_CUDA(cudaMalloc((void **)&mat->d_sblk_ptr, blk_ptr.size() * 4));
            SB_CUDA(cudaMalloc((void **)&mat->d_slen, slen.size() * 4));
            SB_CUDA(cudaMalloc((void **)&mat->d_scol, (size_t)(blocks + 1) * 16));
            SB_CUDA(cudaMalloc((void **)&mat->d_scode, (size_t)(blocks + 1) * 4));
            SB_CUDA(cudaMemcpy(mat->d_row_order, order.data(), order.size() * 4, cudaMemcpyHostToDevice));
            SB_CUDA(cudaMemcpy(mat->d_sblk_ptr, blk_ptr.data(), blk_ptr.size() * 4, cudaMemcpyHostToDevice));
            SB_CUDA(cudaMemcpy(mat->d_slen, slen.data(), slen.size() * 4, cudaMemcpyHostToDevice));
            SB_LAUNCH(sorted_copy_kernel, (mat->n_ordered + 7) / 8, 256, 0, 0, mat->d_row_ptr, mat->d_col, mat->d_code,
                      mat->d_row_order, mat->d_sblk_ptr, mat->n_ordered, reinterpret_cast<int32_t *>(mat->d_scol),
                      reinterpret_cast<uint8_t *>(mat->d_scode));
            SB_CHECK_LAUNCH();
            SB_CUDA(cudaDeviceSynchronize());
            mat->format_bytes += (int64_t)order.size() * 12 + 4 + (blocks + 1) * 20;
        }
    }
    if (mat->n_long) {
        SB_CUDA(cudaMalloc((void **)&mat->d_long_rows, long_rows.size() * 4));
        SB_CUDA(cudaMalloc((void **)&mat->d_long_seg_ptr, seg_ptr.size() * 4));
        SB_CUDA(cudaMalloc((void **)&mat->d_seg_begin, seg_b.size() * 4));
        SB_CUDA(cudaMalloc((void **)&mat->d_seg_end, seg_e.size() * 4));
        SB_CUDA(cudaMemcpy(mat->d_long_rows, long_rows.data(), long_rows.size() * 4, cudaMemcpyHostToDevice));
        SB_CUDA(cudaMemcpy(mat->d_long_seg_ptr, seg_ptr.data(), seg_ptr.size() * 4, cudaMemcpyHostToDevice));
        SB_CUDA(cudaMemcpy(mat->d_seg_begin, seg_b.data(), seg_b.size() * 4, cudaMemcpyHostToDevice));
        SB_CUDA(cudaMemcpy(mat->d_seg_end, seg_e.data(), seg_e.size() * 4, cudaMemcpyHostToDevice));
        mat->format_bytes += (int64_t)(long_rows.size() + seg_ptr.size() + 2 * seg_b.size()) * 4;
    }
    return SBLAS_B200_OK;
}

static int alloc_csr(sblas_b200_matrix *mat, const float *table) {
    const size_t padded = (((size_t)mat->nnz + 3) & ~(size_t)3) + 4;
    SB_CUDA(cudaMalloc((void **)&mat->d_row_ptr, ((size_t)mat->n + 1) * 4));
    SB_CUDA(cudaMalloc((void **)&mat->d_col, padded * 4));
    SB_CUDA(cudaMalloc((void **)&mat->d_code, padded));
    SB_CUDA(cudaMalloc((void **)&mat->d_table, 256 * 4));
    SB_CUDA(cudaMemset(mat->d_col, 0, padded * 4));
    SB_CUDA(cudaMemset(mat->d_code, 0, padded));
    float h_tab[256];
    for (int t = 0; t < 256; ++t) h_tab[t] = (t < mat->T) ? table[t] : 0.0f;
    for (int t = 0; t < mat->T; ++t)
        if (!(h_tab[t] - h_tab[t] == 0.0f)) mat->table_finite = false;      // Inf or NaN in the codebook
    SB_CUDA(cudaMemcpy(mat->d_table, h_tab, sizeof(h_tab), cudaMemcpyHostToDevice));
    mat->format_bytes += ((int64_t)mat->n + 1) * 4 + (int64_t)padded * 5 + 1024;
    return SBLAS_B200_OK;
}

static void release(sblas_b200_matrix *mat) {
    if (!mat) return;
    cudaFree(mat->d_row_ptr); cudaFree(mat->d_col); cudaFree(mat->d_code); cudaFree(mat->d_table);
    cudaFree(mat->d_row_order); cudaFree(mat->d_sblk_ptr); cudaFree(mat->d_slen); cudaFree(mat->d_scol); cudaFree(mat->d_scode);
    cudaFree(mat->d_long_rows); cudaFree(mat->d_long_seg_ptr); cudaFree(mat->d_seg_begin); cudaFree(mat->d_seg_end);
    for (TileFormat &tf : mat->tile) { cudaFree(tf.d_blocks); cudaFree(tf.d_grp_ptr); cudaFree(tf.d_grp_nq); }
    cudaFree(mat->d_ws); cudaFree(mat->d_partial); cudaFree(mat->d_host_a); cudaFree(mat->d_host_c);
    cudaFree(mat->d_host_p); cudaFree(mat->d_dense_code); cudaFree(mat->d_peer_y);
    cudaFree(mat->d_tc_ws);
    cudaFree(mat->native.d_pos); cudaFree(mat->native.d_code); cudaFree(mat->native.d_seg_begin); cudaFree(mat->native.d_seg_end);
    cudaFree(mat->native.d_seg_pos0); cudaFree(mat->native.d_panel_seg_first); cudaFree(mat->native.d_slot_panel);
    cudaFree(mat->native.d_partial);
    cudaFree(mat->blocked.d_lcol); cudaFree(mat->blocked.d_lrow); cudaFree(mat->blocked.d_code);
    cudaFree(mat->blocked.d_seg_ptr); cudaFree(mat->blocked.d_cb_start); cudaFree(mat->blocked.d_prod);
    if (mat->stream) cudaStreamDestroy(mat->stream);
    if (mat->ws_event) cudaEventDestroy(mat->ws_event);
    if (mat->stream2) cudaStreamDestroy(mat->stream2);
    if (mat->stream3) cudaStreamDestroy(mat->stream3);
    if (mat->side_stream) cudaStreamDestroy(mat->side_stream);
    if (mat->side_fork) cudaEventDestroy(mat->side_fork);
    if (mat->side_join) cudaEventDestroy(mat->side_join);
    for (cudaEvent_t ev : mat->host_events) cudaEventDestroy(ev);
    delete mat;
}

// Scratch device arrays of one upload, freed on every exit path.
struct UploadScratch {
    uint8_t *pos = nullptr, *code = nullptr;           // pos: pos_bytes per entry
    int32_t *seg_begin = nullptr, *seg_end = nullptr, *seg_panel = nullptr, *panel_seg_first = nullptr,
            *panel_col_off = nullptr, *seg_hist = nullptr, *row_len = nullptr;
    uint32_t *seg_pos0 = nullptr;
    unsigned long long *fillers = nullptr;
    int32_t *hist_max = nullptr;                       // most entries of one column inside one segment
    ~UploadScratch() {
        cudaFree(pos); cudaFree(code); cudaFree(seg_begin); cudaFree(seg_end); cudaFree(seg_panel);
        cudaFree(panel_seg_first); cudaFree(panel_col_off); cudaFree(seg_hist); cudaFree(row_len);
        cudaFree(seg_pos0); cudaFree(fillers); cudaFree(hist_max);
    }
};

// pos_bytes = sizeof(PosIndex_t): 1 for the reference's instantiation, 2 for SparseMatrix<uint16, uint8, float>
// the largest per-segment column count (seg_hist before panel_bin_scan_kernel turns it into offsets)
__global__ void __launch_bounds__(256)
seg_hist_max_kernel(const int32_t *__restrict__ seg_hist, size_t count, int32_t *__restrict__ out) {
    int32_t m = 0;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < count; i += (size_t)gridDim.x * 256) m = max(m, seg_hist[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}

// Keep the slab's streams and segment table in the handle for src/sparse/spmv_native.cu instead of freeing
// them with the rest of the upload scratch.  One-byte positions only.  Opt-in (SBLAS_B200_KEEP_NATIVE=1): the native kernel
// is measured slower than the CSR kernels everywhere (run_spmv_native), so the 2 bytes per entry are not spent by default.
static int keep_native_streams(sblas_b200_matrix *mat, UploadScratch &sc, int nseg, int np, int pos_bytes,
                               const std::vector<int32_t> &panel_col_off) {
    const char *env = getenv("SBLAS_B200_KEEP_NATIVE");
    const int mode = (env && *env) ? atoi(env) : -1;
    if (mode == 0 || pos_bytes != 1 || nseg == 0 || mat->nnz == 0) return SBLAS_B200_OK;
    int32_t h_max = 0;
    SB_CUDA(cudaMemcpy(&h_max, sc.hist_max, 4, cudaMemcpyDeviceToHost));
    if (mode != 1) return SBLAS_B200_OK;
    NativeFormat &nf = mat->native;
    nf.nseg = nseg;
    nf.npanels = np;
    nf.nslots = (mat->n + kPanelWidth - 1) / kPanelWidth;
    nf.max_col_count = h_max;
    std::vector<int32_t> slot_panel((size_t)nf.nslots, -1);
    for (int p = 0; p < np; ++p) slot_panel[(size_t)(panel_col_off[p] >> kPanelShift)] = p;
    SB_CUDA(cudaMalloc((void **)&nf.d_slot_panel, (size_t)nf.nslots * 4));
    SB_CUDA(cudaMemcpy(nf.d_slot_panel, slot_panel.data(), (size_t)nf.nslots * 4, cudaMemcpyHostToDevice));
    nf.d_pos = sc.pos; sc.pos = nullptr;
    nf.d_code = sc.code; sc.code = nullptr;
    nf.d_seg_begin = sc.seg_begin; sc.seg_begin = nullptr;
    nf.d_seg_end = sc.seg_end; sc.seg_end = nullptr;
    nf.d_seg_pos0 = sc.seg_pos0; sc.seg_pos0 = nullptr;
    nf.d_panel_seg_first = sc.panel_seg_first; sc.panel_seg_first = nullptr;
    nf.state = 1;
    mat->format_bytes += 2 * mat->entries + (int64_t)nseg * 12 + (int64_t)(np + 1 + nf.nslots) * 4;
    return SBLAS_B200_OK;
}

static int upload_stream_impl(const void *pos_index, int pos_bytes, const uint8_t *val_index, int64_t entries,
                              const float *val_table, int32_t T, const int32_t *block_bounds,
                              const int32_t *block_index_bounds, int32_t nblocks, int32_t rows, int32_t cols,
                              int32_t col_begin, int32_t col_end, sblas_b200_matrix *mat) {
    mat->k = rows;
    mat->n = col_end - col_begin;
    mat->col_begin = col_begin;
    mat->T = T;
    // panels of the slab: recorded in ascending col_off (sparse-matrix.cc:35,57-60), so they are contiguous
    int b_first = 0;
    while (b_first < nblocks && block_bounds[2 * b_first + 1] < col_begin) ++b_first;
    int b_last = b_first;
    while (b_last < nblocks && block_bounds[2 * b_last + 1] < col_end) ++b_last;
    const int np = b_last - b_first;
    const int64_t e_begin = np ? block_index_bounds[2 * b_first] : 0;
    const int64_t e_end = np ? block_index_bounds[2 * (b_last - 1) + 1] : 0;
    mat->entries = e_end - e_begin;

    std::vector<int32_t> seg_begin, seg_end, seg_panel, panel_seg_first(np + 1, 0), panel_col_off(np, 0);
    for (int p = 0; p < np; ++p) {
        const int b = b_first + p;
        const int64_t s = block_index_bounds[2 * b], e = block_index_bounds[2 * b + 1];
        if (block_bounds[2 * b] != 0) return fail(SBLAS_B200_EFORMAT, "block %d: row_off %d != 0 (block_row_shift must be 0)", b, block_bounds[2 * b]);
        if (s < 0 || e < s || e > entries) return fail(SBLAS_B200_EFORMAT, "block %d: index bounds [%lld,%lld) outside the stream", b, (long long)s, (long long)e);
        if (p && s != block_index_bounds[2 * (b - 1) + 1]) return fail(SBLAS_B200_EFORMAT, "block %d: index bounds are not contiguous", b);
        if ((block_bounds[2 * b + 1] & kPanelMask) || block_bounds[2 * b + 1] >= cols) return fail(SBLAS_B200_EFORMAT, "block %d: col_off %d is not a panel start inside the matrix", b, block_bounds[2 * b + 1]);
        panel_col_off[p] = block_bounds[2 * b + 1] - col_begin;
        panel_seg_first[p] = (int32_t)seg_begin.size();
        for (int64_t x = s; x < e; x += kSegEntries) {
            seg_begin.push_back((int32_t)(x - e_begin));
            seg_end.push_back((int32_t)(std::min<int64_t>(e, x + kSegEntries) - e_begin));
            seg_panel.push_back(p);
        }
    }
    panel_seg_first[np] = (int32_t)seg_begin.size();
    const int nseg = (int)seg_begin.size();

    UploadScratch sc;
    const int n = mat->n;
    SB_CUDA(cudaMalloc((void **)&sc.row_len, ((size_t)n + 1) * 4));
    SB_CUDA(cudaMemset(sc.row_len, 0, ((size_t)n + 1) * 4));
    SB_CUDA(cudaMalloc((void **)&sc.fillers, 16));
    SB_CUDA(cudaMemset(sc.fillers, 0, 16));
    if (nseg) {
        // (+32 bytes, zeroed: the native-stream kernel reads whole 8-byte words around a segment's last entries)
        SB_CUDA(cudaMalloc((void **)&sc.pos, (size_t)mat->entries * pos_bytes + 32));
        SB_CUDA(cudaMalloc((void **)&sc.code, (size_t)mat->entries + 32));
        SB_CUDA(cudaMemset(sc.pos + (size_t)mat->entries * pos_bytes, 0, 32));
        SB_CUDA(cudaMemset(sc.code + (size_t)mat->entries, 0, 32));
        SB_CUDA(cudaMalloc((void **)&sc.hist_max, 4));
        SB_CUDA(cudaMemset(sc.hist_max, 0, 4));
        SB_CUDA(cudaMemcpy(sc.pos, static_cast<const uint8_t *>(pos_index) + e_begin * pos_bytes, (size_t)mat->entries * pos_bytes,
                           cudaMemcpyDefault));                  // host streams, or the device-resident ones of copyform.cu
        SB_CUDA(cudaMemcpy(sc.code, val_index + e_begin, (size_t)mat->entries, cudaMemcpyDefault));
        SB_CUDA(cudaMalloc((void **)&sc.seg_begin, (size_t)nseg * 4));
        SB_CUDA(cudaMalloc((void **)&sc.seg_end, (size_t)nseg * 4));
        SB_CUDA(cudaMalloc((void **)&sc.seg_panel, (size_t)nseg * 4));
        SB_CUDA(cudaMalloc((void **)&sc.seg_pos0, (size_t)nseg * 4));
        SB_CUDA(cudaMalloc((void **)&sc.seg_hist, (size_t)nseg * kPanelWidth * 4));
        SB_CUDA(cudaMalloc((void **)&sc.panel_seg_first, ((size_t)np + 1) * 4));
        SB_CUDA(cudaMalloc((void **)&sc.panel_col_off, (size_t)np * 4));
        SB_CUDA(cudaMemcpy(sc.seg_begin, seg_begin.data(), (size_t)nseg * 4, cudaMemcpyHostToDevice));
        SB_CUDA(cudaMemcpy(sc.seg_end, seg_end.data(), (size_t)nseg * 4, cudaMemcpyHostToDevice));
        SB_CUDA(cudaMemcpy(sc.seg_panel, seg_panel.data(), (size_t)nseg * 4, cudaMemcpyHostToDevice));
        SB_CUDA(cudaMemcpy(sc.panel_seg_first, panel_seg_first.data(), ((size_t)np + 1) * 4, cudaMemcpyHostToDevice));
        SB_CUDA(cudaMemcpy(sc.panel_col_off, panel_col_off.data(), (size_t)np * 4, cudaMemcpyHostToDevice));

        const int seg_grid = (nseg + kWarpsPerCta - 1) / kWarpsPerCta;
        const uint16_t *pos16 = reinterpret_cast<const uint16_t *>(sc.pos);
        if (pos_bytes == 2) SB_LAUNCH(seg_delta_sum_kernel<uint16_t>, seg_grid, 32 * kWarpsPerCta, 0, 0, pos16, sc.seg_begin, sc.seg_end, nseg, sc.seg_pos0);
        else SB_LAUNCH(seg_delta_sum_kernel<uint8_t>, seg_grid, 32 * kWarpsPerCta, 0, 0, sc.pos, sc.seg_begin, sc.seg_end, nseg, sc.seg_pos0);
        SB_LAUNCH(panel_seg_scan_kernel, (np + 127) / 128, 128, 0, 0, sc.panel_seg_first, np, sc.seg_pos0);
        if (pos_bytes == 2) SB_LAUNCH(seg_count_kernel<uint16_t>, seg_grid, 32 * kWarpsPerCta, 0, 0, pos16, sc.code, T, sc.seg_begin, sc.seg_end,
                                      sc.seg_panel, sc.panel_col_off, sc.seg_pos0, nseg, n, mat->k, sc.seg_hist, sc.fillers);
        else SB_LAUNCH(seg_count_kernel<uint8_t>, seg_grid, 32 * kWarpsPerCta, 0, 0, sc.pos, sc.code, T, sc.seg_begin, sc.seg_end,
                       sc.seg_panel, sc.panel_col_off, sc.seg_pos0, nseg, n, mat->k, sc.seg_hist, sc.fillers);
        SB_LAUNCH(seg_hist_max_kernel, 148 * 4, 256, 0, 0, sc.seg_hist, (size_t)nseg * kPanelWidth, sc.hist_max);
        SB_LAUNCH(panel_bin_scan_kernel, np, kPanelWidth, 0, 0, sc.panel_seg_first, sc.panel_col_off, sc.seg_hist,
                  sc.row_len, n);
        SB_CHECK_LAUNCH();
    }
    // row_ptr: needs nnz first to size col/code -> scan into a temporary, then allocate
    int32_t *d_row_ptr_tmp = nullptr;
    SB_CUDA(cudaMalloc((void **)&d_row_ptr_tmp, ((size_t)n + 1) * 4));
    SB_LAUNCH(row_ptr_scan_kernel, 1, 1024, 0, 0, sc.row_len, n, d_row_ptr_tmp);
    std::vector<int32_t> h_row_ptr((size_t)n + 1);
    cudaError_t e = cudaMemcpy(h_row_ptr.data(), d_row_ptr_tmp, ((size_t)n + 1) * 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { cudaFree(d_row_ptr_tmp); return cuda_fail(e, "row_ptr readback", __FILE__, __LINE__); }
    unsigned long long h_fill[2] = {0, 0};
    e = cudaMemcpy(h_fill, sc.fillers, 16, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { cudaFree(d_row_ptr_tmp); return cuda_fail(e, "filler readback", __FILE__, __LINE__); }
    mat->nnz = h_row_ptr[n];
    mat->fillers = (int64_t)h_fill[0];
    if (h_fill[1]) {
        cudaFree(d_row_ptr_tmp);
        return fail(SBLAS_B200_EFORMAT, "%llu stream entries decode to positions outside the %d x %d matrix", h_fill[1], mat->k, n);
    }
    if (mat->nnz + mat->fillers != mat->entries) {
        cudaFree(d_row_ptr_tmp);
        return fail(SBLAS_B200_EFORMAT, "stream accounting: nnz %lld + fillers %lld != entries %lld", (long long)mat->nnz,
                    (long long)mat->fillers, (long long)mat->entries);
    }
    int rc = alloc_csr(mat, val_table);
    if (rc != SBLAS_B200_OK) { cudaFree(d_row_ptr_tmp); return rc; }
    e = cudaMemcpy(mat->d_row_ptr, d_row_ptr_tmp, ((size_t)n + 1) * 4, cudaMemcpyDeviceToDevice);
    cudaFree(d_row_ptr_tmp);
    if (e != cudaSuccess) return cuda_fail(e, "row_ptr copy", __FILE__, __LINE__);
    if (nseg) {
        const int seg_grid = (nseg + kWarpsPerCta - 1) / kWarpsPerCta;
        if (pos_bytes == 2)
            SB_LAUNCH(seg_scatter_kernel<uint16_t>, seg_grid, 32 * kWarpsPerCta, 0, 0, reinterpret_cast<const uint16_t *>(sc.pos), sc.code, T,
                      sc.seg_begin, sc.seg_end, sc.seg_panel, sc.seg_pos0, nseg, sc.panel_col_off, sc.seg_hist, mat->d_row_ptr, n, mat->k,
                      mat->d_col, mat->d_code);
        else
            SB_LAUNCH(seg_scatter_kernel<uint8_t>, seg_grid, 32 * kWarpsPerCta, 0, 0, sc.pos, sc.code, T, sc.seg_begin, sc.seg_end,
                      sc.seg_panel, sc.seg_pos0, nseg, sc.panel_col_off, sc.seg_hist, mat->d_row_ptr, n, mat->k,
                      mat->d_col, mat->d_code);
        SB_CHECK_LAUNCH();
    }
    SB_CUDA(cudaDeviceSynchronize());
    rc = plan_spmv(mat, h_row_ptr.data());
    if (rc != SBLAS_B200_OK) return rc;
    return keep_native_streams(mat, sc, nseg, np, pos_bytes, panel_col_off);
}

static int check_stream_args(const void *pos_index, const uint8_t *val_index, int64_t entries, const float *val_table,
                             int32_t T, const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                             int32_t rows, int32_t cols, sblas_b200_matrix **out) {
    if (!out) return fail(SBLAS_B200_EINVAL, "out is null");
    *out = nullptr;
    if (T < 0 || T > 255) return fail(SBLAS_B200_EINVAL, "valid_table_size %d outside [0,255]", T);   // sparse-matrix.cc:25
    if (entries < 0 || nblocks < 0 || rows < 0 || cols < 0) return fail(SBLAS_B200_EINVAL, "negative size");
    if (entries >= ((int64_t)1 << 31)) return fail(SBLAS_B200_EINVAL, "entries do not fit int32");
    if ((int64_t)rows * kPanelWidth > ((int64_t)1 << 31)) return fail(SBLAS_B200_EINVAL, "rows*256 exceeds the reference's int position range");
    if (entries && (!pos_index || !val_index)) return fail(SBLAS_B200_EINVAL, "null stream pointer");
    if (T && !val_table) return fail(SBLAS_B200_EINVAL, "null val_table");
    if (nblocks && (!block_bounds || !block_index_bounds)) return fail(SBLAS_B200_EINVAL, "null block arrays");
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200

using namespace sblas_b200;

extern "C" {

int32_t sblas_b200_abi_version(void) { return SBLAS_B200_ABI_VERSION; }
const char *sblas_b200_last_error(void) { return sblas_b200::t_error; }
int64_t sblas_b200_launch_count(void) { return g_kernel_launches.load(); }

int sblas_b200_device_count(int32_t *count) {
    if (!count) return fail(SBLAS_B200_EINVAL, "count is null");
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess) { *count = 0; return cuda_fail(e, "cudaGetDeviceCount", __FILE__, __LINE__); }
    *count = c;
    return SBLAS_B200_OK;
}

static int upload_slab_any(const void *pos_index, int pos_bytes, const uint8_t *val_index, int64_t entries, const float *val_table,
                           int32_t T, const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                           int32_t rows, int32_t cols, int32_t col_begin, int32_t col_end, int32_t device,
                           sblas_b200_matrix **out) {
    int rc = check_stream_args(pos_index, val_index, entries, val_table, T, block_bounds, block_index_bounds, nblocks, rows, cols, out);
    if (rc != SBLAS_B200_OK) return rc;
    if (col_begin < 0 || col_end < col_begin || col_end > cols) return fail(SBLAS_B200_EINVAL, "slab [%d,%d) outside [0,%d)", col_begin, col_end, cols);
    if (col_begin & kPanelMask) return fail(SBLAS_B200_EINVAL, "col_begin %d is not panel aligned (multiple of 256)", col_begin);
    if ((col_end & kPanelMask) && col_end != cols) return fail(SBLAS_B200_EINVAL, "col_end %d is neither panel aligned nor the matrix edge", col_end);
    DeviceScope scope(device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    sblas_b200_matrix *mat = new (std::nothrow) sblas_b200_matrix();
    if (!mat) return fail(SBLAS_B200_ENOMEM, "host allocation failed");
    mat->device = device;
    rc = upload_stream_impl(pos_index, pos_bytes, val_index, entries, val_table, T, block_bounds, block_index_bounds, nblocks, rows,
                            cols, col_begin, col_end, mat);
    if (rc != SBLAS_B200_OK) { release(mat); return rc; }
    *out = mat;
    return SBLAS_B200_OK;
}

int sblas_b200_upload_slab(const uint8_t *pos_index, const uint8_t *val_index, int64_t entries, const float *val_table,
                           int32_t T, const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                           int32_t rows, int32_t cols, int32_t col_begin, int32_t col_end, int32_t device,
                           sblas_b200_matrix **out) {
    return upload_slab_any(pos_index, 1, val_index, entries, val_table, T, block_bounds, block_index_bounds, nblocks, rows, cols,
                           col_begin, col_end, device, out);
}

int sblas_b200_upload_slab16(const uint16_t *pos_index, const uint8_t *val_index, int64_t entries, const float *val_table,
                             int32_t T, const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                             int32_t rows, int32_t cols, int32_t col_begin, int32_t col_end, int32_t device,
                             sblas_b200_matrix **out) {
    return upload_slab_any(pos_index, 2, val_index, entries, val_table, T, block_bounds, block_index_bounds, nblocks, rows, cols,
                           col_begin, col_end, device, out);
}

int sblas_b200_upload(const uint8_t *pos_index, const uint8_t *val_index, int64_t entries, const float *val_table,
                      int32_t T, const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks,
                      int32_t rows, int32_t cols, int32_t device, sblas_b200_matrix **out) {
    return sblas_b200_upload_slab(pos_index, val_index, entries, val_table, T, block_bounds, block_index_bounds, nblocks,
                                  rows, cols, 0, cols, device, out);
}

int sblas_b200_upload_csr(const int32_t *row_ptr, const int32_t *col, const uint8_t *code, const float *table, int32_t T,
                          int32_t n, int32_t k, int32_t device, sblas_b200_matrix **out) {
    if (!out) return fail(SBLAS_B200_EINVAL, "out is null");
    *out = nullptr;
    if (T < 0 || T > 255) return fail(SBLAS_B200_EINVAL, "valid_table_size %d outside [0,255]", T);
    if (n < 0 || k < 0 || !row_ptr) return fail(SBLAS_B200_EINVAL, "bad CSR shape");
    if (row_ptr[0] != 0) return fail(SBLAS_B200_EFORMAT, "row_ptr[0] != 0");
    for (int j = 0; j < n; ++j)
        if (row_ptr[j + 1] < row_ptr[j]) return fail(SBLAS_B200_EFORMAT, "row_ptr not monotone at row %d", j);
    const int64_t nnz = row_ptr[n];
    if (nnz && (!col || !code || !table)) return fail(SBLAS_B200_EINVAL, "null CSR array");
    DeviceScope scope(device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    sblas_b200_matrix *mat = new (std::nothrow) sblas_b200_matrix();
    if (!mat) return fail(SBLAS_B200_ENOMEM, "host allocation failed");
    mat->device = device; mat->k = k; mat->n = n; mat->T = T; mat->nnz = nnz; mat->entries = nnz;
    int rc = alloc_csr(mat, table);
    if (rc == SBLAS_B200_OK) {
        cudaError_t e = cudaMemcpy(mat->d_row_ptr, row_ptr, ((size_t)n + 1) * 4, cudaMemcpyHostToDevice);
        if (e == cudaSuccess && nnz) e = cudaMemcpy(mat->d_col, col, (size_t)nnz * 4, cudaMemcpyHostToDevice);
        if (e == cudaSuccess && nnz) e = cudaMemcpy(mat->d_code, code, (size_t)nnz, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) rc = cuda_fail(e, "CSR upload", __FILE__, __LINE__);
    }
    if (rc == SBLAS_B200_OK && nnz) {
        int *d_bad = nullptr, h_bad = 0;
        cudaError_t e = cudaMalloc((void **)&d_bad, 4);
        if (e == cudaSuccess) e = cudaMemset(d_bad, 0, 4);
        if (e == cudaSuccess) {
            SB_LAUNCH(csr_validate_kernel, 148 * 8, 256, 0, 0, mat->d_col, mat->d_code, nnz, k, T, d_bad);
            SB_LAUNCH(csr_sorted_kernel, (n + 7) / 8, 256, 0, 0, mat->d_row_ptr, mat->d_col, n, d_bad);
            e = cudaMemcpy(&h_bad, d_bad, 4, cudaMemcpyDeviceToHost);
        }
        cudaFree(d_bad);
        if (e != cudaSuccess) rc = cuda_fail(e, "CSR validation", __FILE__, __LINE__);
        else if (h_bad & 2) mat->csr_sorted = false;
        if (e == cudaSuccess && (h_bad & 1)) rc = fail(SBLAS_B200_EFORMAT, "CSR holds a column index outside [0,%d) or a code >= %d", k, T);
    }
    if (rc == SBLAS_B200_OK) rc = plan_spmv(mat, row_ptr);
    if (rc != SBLAS_B200_OK) { release(mat); return rc; }
    *out = mat;
    return SBLAS_B200_OK;
}

int sblas_b200_partition(const int32_t *block_bounds, const int32_t *block_index_bounds, int32_t nblocks, int32_t cols,
                         int32_t parts, int32_t *col_splits) {
    if (parts < 1 || cols < 0 || !col_splits || nblocks < 0) return fail(SBLAS_B200_EINVAL, "bad partition arguments");
    if (nblocks && (!block_bounds || !block_index_bounds)) return fail(SBLAS_B200_EINVAL, "null block arrays");
    const int npanels = (cols + kPanelWidth - 1) / kPanelWidth;
    std::vector<int64_t> prefix((size_t)npanels + 1, 0);         // entries in panels [0, p)
    for (int b = 0; b < nblocks; ++b) {
        const int p = block_bounds[2 * b + 1] >> kPanelShift;
        if (p < 0 || p >= npanels) return fail(SBLAS_B200_EFORMAT, "block %d: col_off outside the matrix", b);
        prefix[p + 1] = (int64_t)block_index_bounds[2 * b + 1] - block_index_bounds[2 * b];
    }
    for (int p = 0; p < npanels; ++p) prefix[p + 1] += prefix[p];
    const int64_t total = prefix[npanels];
    col_splits[0] = 0;
    int prev = 0;
    for (int g = 1; g < parts; ++g) {
        // panel boundary whose prefix is closest to the ideal share, never moving backwards
        const double target = (double)total * g / parts;
        int b = (int)(std::lower_bound(prefix.begin(), prefix.end(), (int64_t)(target + 0.5)) - prefix.begin());
        if (b > npanels) b = npanels;
        if (b > 0 && target - (double)prefix[b - 1] < (double)prefix[b] - target) --b;
        if (b < prev) b = prev;
        prev = b;
        col_splits[g] = std::min(cols, b * kPanelWidth);
    }
    col_splits[parts] = cols;
    return SBLAS_B200_OK;
}

int sblas_b200_partition_csr(const int32_t *row_ptr, int32_t n, int32_t parts, int32_t align, int32_t *row_splits) {
    if (!row_ptr || n < 0 || parts < 1 || align < 1 || (align & (align - 1)) || !row_splits)
        return fail(SBLAS_B200_EINVAL, "bad partition arguments");
    const int64_t total = row_ptr[n];
    row_splits[0] = 0;
    int prev = 0;
    for (int g = 1; g < parts; ++g) {
        // aligned row whose prefix is closest to the ideal share, never moving backwards
        const int64_t target = (int64_t)((double)total * g / parts + 0.5);
        int r = (int)(std::lower_bound(row_ptr, row_ptr + n + 1, (int32_t)std::min<int64_t>(target, total)) - row_ptr);
        int lo = r / align * align, hi = std::min<int64_t>(n, (int64_t)lo + align);
        if (hi == n && (n % align)) hi = lo;                  // interior splits stay aligned
        r = (target - row_ptr[lo] <= (int64_t)row_ptr[hi] - target) ? lo : hi;
        if (r < prev) r = prev;
        prev = r;
        row_splits[g] = r;
    }
    row_splits[parts] = n;
    return SBLAS_B200_OK;
}

int sblas_b200_free(sblas_b200_matrix *mat) {
    if (!mat) return SBLAS_B200_OK;
    DeviceScope scope(mat->device);
    forget_host_buffers(mat, nullptr);                      // caller buffers this handle's host calls had registered
    release(mat);
    return SBLAS_B200_OK;
}

int sblas_b200_get_info(const sblas_b200_matrix *mat, sblas_b200_info *info) {
    if (!mat || !info) return fail(SBLAS_B200_EINVAL, "null argument");
    memset(info, 0, sizeof(*info));
    info->device = mat->device;
    info->s_rows = mat->k;
    info->s_cols = mat->n;
    info->col_begin = mat->col_begin;
    info->table_size = mat->T;
    info->max_row_nnz = mat->max_row_nnz;
    info->long_rows = mat->n_long;
    info->spmv_lanes = mat->spmv_lanes;
    info->nnz = mat->nnz;
    info->fillers = mat->fillers;
    info->entries = mat->entries;
    info->device_bytes = mat->format_bytes + (int64_t)(mat->ws_bytes + mat->partial_bytes + mat->host_a_bytes + mat->host_c_bytes);
    memcpy(info->row_len_hist, mat->row_len_hist, sizeof(info->row_len_hist));
    return SBLAS_B200_OK;
}

int sblas_b200_decode_dense(const sblas_b200_matrix *mat, float *dense, int32_t stride, int32_t trans) {
    if (!mat || (!dense && mat->n && mat->k)) return fail(SBLAS_B200_EINVAL, "null argument");
    const int64_t lead = trans ? mat->n : mat->k;
    const int64_t minor = trans ? mat->k : mat->n;
    if (stride < minor) return fail(SBLAS_B200_EINVAL, "stride %d < %lld", stride, (long long)minor);
    if (lead == 0 || stride == 0) return SBLAS_B200_OK;
    DeviceScope scope(mat->device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    const size_t bytes = (size_t)lead * stride * 4;
    cudaPointerAttributes attr;
    const bool on_device = cudaPointerGetAttributes(&attr, dense) == cudaSuccess && attr.type == cudaMemoryTypeDevice;
    cudaGetLastError();
    float *d_dense = on_device ? dense : nullptr;                  // device caller: decode in place, no staging
    if (!on_device) SB_CUDA(cudaMalloc((void **)&d_dense, bytes));
    cudaError_t e = cudaMemset(d_dense, 0, bytes);                 // memset(density_matrix, 0, ...) sparse-matrix.cc:109,123
    if (e == cudaSuccess && mat->n) {
        SB_LAUNCH(decode_dense_kernel, (mat->n + 7) / 8, 256, 0, 0, mat->d_row_ptr, mat->d_col, mat->d_code, mat->d_table,
                  mat->n, d_dense, stride, trans);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = on_device ? cudaDeviceSynchronize() : cudaMemcpy(dense, d_dense, bytes, cudaMemcpyDeviceToHost);
    if (!on_device) cudaFree(d_dense);
    if (e != cudaSuccess) return cuda_fail(e, "decode_dense", __FILE__, __LINE__);
    return SBLAS_B200_OK;
}

int sblas_b200_export_csr(const sblas_b200_matrix *mat, int32_t *row_ptr, int32_t *col, uint8_t *code) {
    if (!mat || !row_ptr) return fail(SBLAS_B200_EINVAL, "null argument");
    DeviceScope scope(mat->device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    SB_CUDA(cudaMemcpy(row_ptr, mat->d_row_ptr, ((size_t)mat->n + 1) * 4, cudaMemcpyDeviceToHost));
    if (mat->nnz) {
        if (!col || !code) return fail(SBLAS_B200_EINVAL, "null argument");
        SB_CUDA(cudaMemcpy(col, mat->d_col, (size_t)mat->nnz * 4, cudaMemcpyDeviceToHost));
        SB_CUDA(cudaMemcpy(code, mat->d_code, (size_t)mat->nnz, cudaMemcpyDeviceToHost));
    }
    return SBLAS_B200_OK;
}

}  // extern "C"
