// spmv_native.cu -- K-spmv (native-stream form): y = beta*y + alpha * W x multiplied straight from the reference's two
// byte streams, pos_index_ / val_index_, as CopyForm wrote them (sparse-matrix.cc:21-99) and as the reference's own loop
// walks them (kernel.cc:780-796): 2 bytes per stream entry instead of 5 per non-zero of the CSR copy, and -- the point
// for large matrices -- no gather: a panel's stream enumerates its entries row by row of S, so x is read front to back
// while the 256 outputs of the panel are the scattered side, and those live in shared memory.
//
//   segments  the upload cuts every panel's stream into segments of 2048 entries and knows the linear position
//             (row * 256 + column in panel) each one starts from (device_format.cu: the same segments its conversion
//             kernels use).  One CTA per segment.
//   decode    a warp takes 256 consecutive entries, 32 per step (lane = entry): warp scan of the one-byte deltas + carry ->
//             positions.  Fillers (code >= T, sparse-matrix.cc:46-51) only move the position.
//   products  term = (val_table[code] * alpha) * x[row]   (kernel.cc:791: alpha folded into the weight, multiply, then add)
//   sums      into the warp's own copy of the panel's 256 outputs in shared memory, the rows of S inside a step one after the
//             other (no two lanes on one word), so an output's terms are added in the order the reference adds them --
//             deterministic; the eight warps' sums are added in warp order.
//   finish    spmv_native_finish_kernel adds a panel's segment sums in segment order and applies beta; columns without
//             entries only get the scale (row_ptr of the CSR copy, like every other kernel here).
//
// Absent entries never touch x (only real entries load it), as in the reference.
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "device_matrix.h"

namespace sblas_b200 {

namespace {

constexpr int kNatSeg = 2048;            // stream entries per segment (device_format.cu kSegEntries)
constexpr int kNatThreads = 256;         // = columns of a panel

__device__ __forceinline__ void red_add_f32(float *p, float v) { asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory"); }

// One CTA per segment, one warp per 256 consecutive entries of it, 32 consecutive entries per step (lane = entry).
// The running position goes through the warp as in the upload's own kernels (warp scan of the deltas + carry); a warp's
// starting position is the segment's plus the delta sums of the warps before it (first pass, 8 bytes per lane).
// Terms are added into the warp's private copy of the panel's 256 outputs in shared memory.  Two lanes of a step name the
// same column only when the step spans several rows of S: the rows of a step are added one after the other (REDUX.MIN over
// the lanes still pending), within a row every column is different, so an output's terms are added in stream order.  Thread j finally adds the eight warps'
// sums of column j in warp order.  (First version: entries filed per column through shared-memory counters and put into
// stream order by a per-thread insertion sort -- 5 to 7 times slower, profiles/r02_spmv_native.md.)
__global__ void __launch_bounds__(kNatThreads)
spmv_native_segment_kernel(const uint8_t *__restrict__ pos, const uint8_t *__restrict__ code, const int32_t *__restrict__ seg_begin,
                           const int32_t *__restrict__ seg_end, const uint32_t *__restrict__ seg_pos0,
                           const float *__restrict__ table, int T, float alpha, const float *__restrict__ x,
                           float *__restrict__ partial) {
    constexpr int kWarps = kNatThreads / 32, kPerWarp = kNatSeg / kWarps;      // 8 warps x 256 entries
    __shared__ float tab_s[256];
    __shared__ float y_w[kWarps][256];
    __shared__ uint32_t wsum[kWarps];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, seg = blockIdx.x;
    tab_s[t] = t < T ? table[t] * alpha : 0.0f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) y_w[w][t] = 0.0f;
    const int b = seg_begin[seg] + warp * kPerWarp, e = seg_end[seg];
    // delta sum of this warp's entries
    uint32_t mine = 0;
#pragma unroll
    for (int s = 0; s < kPerWarp / 32; ++s) {
        const int i = b + 32 * s + lane;
        if (i < e) mine += pos[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if (lane == 0) wsum[warp] = mine;
    __syncthreads();
    uint32_t running = seg_pos0[seg];
    for (int w = 0; w < warp; ++w) running += wsum[w];
    float *y_mine = y_w[warp];
    for (int s = 0; s < kPerWarp / 32; ++s) {
        const int i = b + 32 * s;
        if (i >= e) break;                                       // warp-uniform
        const bool in = i + lane < e;
        const uint32_t d = in ? pos[i + lane] : 0u;
        const int c = in ? code[i + lane] : 255;
        uint32_t incl = d;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t p = running + incl;                      // pos_offset += ppos[i]   (kernel.cc:781)
        running += __shfl_sync(0xffffffffu, incl, 31);
        const bool real = in && c < T;                          // code >= T: a filler, it only carried the position on
        const int col = (int)(p & 255u);
        const float term = real ? tab_s[c] * __ldg(x + (p >> 8)) : 0.0f;
        // positions of a step increase strictly: when they span fewer than 256, their low bytes -- the columns -- all differ
        const uint32_t span = __shfl_sync(0xffffffffu, incl, 31) - __shfl_sync(0xffffffffu, incl, 0);
        if (span < 256u) {
            if (real) y_mine[col] += term;
            __syncwarp();
            continue;
        }
        // otherwise the rows of S one after the other (within a row every column is different): stream order, no two lanes
        // on one word
        const unsigned row = real ? (p >> 8) : 0xffffffffu;
        bool pending = real;
        for (unsigned todo = __ballot_sync(0xffffffffu, pending); todo; todo = __ballot_sync(0xffffffffu, pending)) {
            const unsigned cur = __reduce_min_sync(0xffffffffu, pending ? row : 0xffffffffu);
            if (pending && row == cur) { y_mine[col] += term; pending = false; }
            __syncwarp();
        }
    }
    __syncthreads();
    float acc = y_w[0][t];
#pragma unroll
    for (int w = 1; w < kWarps; ++w) acc += y_w[w][t];
    partial[(size_t)seg * 256 + t] = acc;
}

// grid = slots of 256 columns; 4 x 256 threads: thread (q, column) adds the q-th quarter of the panel's segment sums in
// segment order, the four quarters are then added in order (a fixed association: deterministic), then beta.
constexpr int kFinQ = 4;
__global__ void __launch_bounds__(kFinQ * 256)
spmv_native_finish_kernel(const float *__restrict__ partial, const int32_t *__restrict__ slot_panel,
                          const int32_t *__restrict__ panel_seg_first, const int32_t *__restrict__ row_ptr, int n,
                          float *__restrict__ y, float beta, int red) {
    __shared__ float quarter[kFinQ][256];
    const int jj = threadIdx.x & 255, q = threadIdx.x >> 8, j = blockIdx.x * 256 + jj;
    const int p = slot_panel[blockIdx.x];
    float sum = 0.0f;
    if (p >= 0 && j < n) {
        const int f = panel_seg_first[p], cnt = panel_seg_first[p + 1] - f;
        const int s0 = f + (int)((int64_t)cnt * q / kFinQ), s1 = f + (int)((int64_t)cnt * (q + 1) / kFinQ);
        int sg = s0;
        for (; sg + 8 <= s1; sg += 8) {                         // eight loads in flight, added in order
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = partial[(size_t)(sg + u) * 256 + jj];
#pragma unroll
            for (int u = 0; u < 8; ++u) sum += v[u];
        }
        for (; sg < s1; ++sg) sum += partial[(size_t)sg * 256 + jj];
    }
    quarter[q][jj] = sum;
    __syncthreads();
    if (q != 0 || j >= n) return;
#pragma unroll
    for (int u = 1; u < kFinQ; ++u) sum += quarter[u][jj];
    const int len = row_ptr[j + 1] - row_ptr[j];
    if (red) {                                                  // y in a peer GPU's memory, beta == 1: add at the owner
        if (len) red_add_f32(y + j, sum);
        return;
    }
    if (len == 0 && beta == 1.0f) return;                       // column without entries, no scale: untouched
    float c = y[j];
    if (beta != 1.0f) c *= beta;
    y[j] = len ? c + sum : c;                                   // no entry: exactly the scaled value (keeps -0)
}

}  // namespace

// SBLAS_B200_SPMV_NATIVE = 0 never | 1 whenever the streams were kept (read per call: tests flip it).
int run_spmv_native(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta, cudaStream_t stream) {
    NativeFormat &nf = mat->native;
    if (nf.state != 1) return 1;
    const char *env = getenv("SBLAS_B200_SPMV_NATIVE");
    const int mode = (env && *env) ? atoi(env) : -1;
    if (mode == 0) return 1;
    // Not chosen on its own: on every shape measured the CSR row kernels are 1.5-3 times faster -- their x gathers hit L1 /
    // L2 up to k of a few 10^5, and they spend about one instruction per entry where this one spends a warp scan and an
    // ordered shared-memory update (profiles/r02_spmv_native.md).  It is there for callers that cannot afford the 5 B per
    // non-zero of the CSR copy next to the streams (SBLAS_B200_KEEP_NATIVE=1 at upload, SBLAS_B200_SPMV_NATIVE=1 per call).
    if (mode != 1) return 1;
    const bool red = beta == 1.0f && is_peer_pointer(y);
    if (nf.nseg && !nf.d_partial) {
        if (cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone; cudaStreamIsCapturing(stream, &cap) == cudaSuccess && cap != cudaStreamCaptureStatusNone)
            return 1;                                           // no allocation inside a capture
        cudaGetLastError();
        if (cudaMalloc((void **)&nf.d_partial, (size_t)nf.nseg * 256 * 4) != cudaSuccess) {   // no room: the other kernels
            cudaGetLastError();
            nf.d_partial = nullptr;
            return 1;
        }
    }
    struct Timer {                                              // event pair around both kernels when profiling
        sblas_b200_matrix *mat; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
        Timer(sblas_b200_matrix *m_, cudaStream_t s_) : mat(m_), st(s_) {
            if (mat->profiling) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, st); }
        }
        ~Timer() { if (a) { cudaEventRecord(b, st); mat->prof_events.push_back(std::make_pair(a, b)); } }
    } timer(mat, stream);
    // 2 bytes per stream entry + the segment sums written and read once (1 KB per 2048 entries)
    note_product(mat, red ? "spmv_native red" : "spmv_native", 1,
                 (2.0 * (double)mat->entries + 8.0 * 256.0 * (double)nf.nseg) / (double)std::max<int64_t>(1, mat->nnz),
                 (int64_t)12 * nf.nseg + (int64_t)4 * (mat->n + 1));
    if (nf.nseg)
        SB_LAUNCH(spmv_native_segment_kernel, nf.nseg, kNatThreads, 0, stream, nf.d_pos, nf.d_code, nf.d_seg_begin, nf.d_seg_end,
                  nf.d_seg_pos0, mat->d_table, mat->T, alpha, x, nf.d_partial);
    if (nf.nslots)
        SB_LAUNCH(spmv_native_finish_kernel, nf.nslots, kFinQ * 256, 0, stream, nf.d_partial, nf.d_slot_panel, nf.d_panel_seg_first,
                  mat->d_row_ptr, mat->n, y, beta, red ? 1 : 0);
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200
