// spmv_blocked.cu -- K-spmv-blocked: SpMV (m == 1) in two streaming passes over the column-blocked
// copy of W (src/utils/blocked_format.cu), for matrices whose x gathers would each cost an L2 sector.
//
// Replaces, like the row kernels of product.cu, the reference's decode + AXPY pair at m = 1
// (kernel.cc:771-800 with the scalar tail :568-582): y[j] = beta*y[j] + sum_r x[r] * fl(table[code]*alpha).
//
//   pass 1  persistent CTAs walk the slot stream; the x tile of the current column block (cbw floats)
//           sits in shared memory, the codebook * alpha is replicated per lane (conflict-free look-up):
//           p[slot] = fl(x_s[lcol] * tab[code])          3 B read + 4 B written per entry, no global gather
//   pass 2  one warp per row sub-block (rs rows, sums in shared memory) walks its sub-segments in column
//           block order -- the entries of one row arrive in column order, the order of the reference's
//           accumulation -- four neighbouring slots per lane (8 + 16 byte loads), and adds each run of equal
//           rows to the row's sum: 6 B read per entry.  No CTA barrier anywhere: the warps of a CTA own
//           disjoint rows.
// Products and sums are separately rounded like the reference's generic build (no FMA).  A run that spans
// lanes is added in phases, lane after lane (fixed order: results are run-to-run identical).
// HBM traffic: 13 B per entry + x + y, against 5 B per entry and one L2 sector per entry for the row kernels.
#include <stdlib.h>

#include <algorithm>

#include "device_matrix.h"

namespace sblas_b200 {

namespace {

constexpr int kP1Threads = 1024;
constexpr int kP2Warps = 16;

template <int U>
__global__ void __launch_bounds__(kP1Threads, 1)
spmv_blocked_pass1_kernel(const uint2 *__restrict__ lcol4, const uint32_t *__restrict__ code4,
                          const int32_t *__restrict__ cb_start, int ncb, int cbw, const float *__restrict__ x, int k,
                          const float *__restrict__ table, int T, float alpha, float4 *__restrict__ prod4, int x_vec) {
    extern __shared__ __align__(16) float smem[];
    float *tab_s = smem;                                    // [256][32]: entry c of lane l at c * 32 + l
    float *x_s = smem + 256 * 32;                           // [cbw]
    const int lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 256 * 32; i += kP1Threads) {
        const int c = i >> 5;
        tab_s[i] = c < T ? table[c] * alpha : 0.0f;         // val_table[code] * alpha (kernel.cc:791)
    }
    const int total_q = cb_start[ncb] >> 2;                 // quads of slots
    constexpr int kStep = kP1Threads * U;
    const int q_first = (int)(((int64_t)total_q * blockIdx.x / gridDim.x) / kStep * kStep);
    const int q_end = blockIdx.x + 1 == gridDim.x ? total_q
                                                  : (int)(((int64_t)total_q * (blockIdx.x + 1) / gridDim.x) / kStep * kStep);
    int q = q_first, cb = 0;
    const float *tab_l = tab_s + lane;
    while (q < q_end) {
        while ((cb_start[cb + 1] >> 2) <= q) ++cb;          // the block slot q lies in (empty blocks skipped)
        const int stop = min(q_end, cb_start[cb + 1] >> 2);
        __syncthreads();                                    // everyone is done with the previous tile (and the table is in)
        const int c0 = cb * cbw, cols = min(cbw, k - c0);
        if (x_vec) {
            const float4 *xg = reinterpret_cast<const float4 *>(x + c0);
            for (int i = threadIdx.x; i < (cols >> 2); i += kP1Threads) reinterpret_cast<float4 *>(x_s)[i] = xg[i];
            for (int i = (cols & ~3) + threadIdx.x; i < cols; i += kP1Threads) x_s[i] = x[c0 + i];
        } else {
            for (int i = threadIdx.x; i < cols; i += kP1Threads) x_s[i] = x[c0 + i];
        }
        __syncthreads();
        for (int qq = q + threadIdx.x; qq < stop; qq += kStep) {
            uint2 lc[U];
            uint32_t cd[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (qq + u * kP1Threads < stop) { lc[u] = __ldcs(lcol4 + qq + u * kP1Threads); cd[u] = __ldcs(code4 + qq + u * kP1Threads); }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (qq + u * kP1Threads < stop) {
                    float4 p;
                    p.x = x_s[lc[u].x & 0xffffu] * tab_l[(cd[u] << 5) & 0x1fe0u];
                    p.y = x_s[lc[u].x >> 16] * tab_l[(cd[u] >> 3) & 0x1fe0u];
                    p.z = x_s[lc[u].y & 0xffffu] * tab_l[(cd[u] >> 11) & 0x1fe0u];
                    p.w = x_s[lc[u].y >> 16] * tab_l[(cd[u] >> 19) & 0x1fe0u];
                    __stcs(prod4 + qq + u * kP1Threads, p);
                }
            }
        }
        q = stop;
    }
}

__device__ __forceinline__ void red_add_f32(float *p, float v) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

// One step of a warp is a chunk of 32 quads (128 slots): lane l holds four neighbouring slots, i.e. rows r0 <= r1 <= r2 <= r3
// (0xffff = padding) and their products.  Inside the lane the products of equal rows are added left to right (column order).
// Only one kind of run crosses lanes: a lane's first run may continue the previous lane's last one.  Such a head run is held
// back and added in a later phase -- phase d for a lane d lanes away from where the run began -- so that no two lanes ever
// update the same row at once and every sum is formed in a fixed order.  D = chunks a warp keeps in flight.
template <int D>
__global__ void __launch_bounds__(32 * kP2Warps, 2)
spmv_blocked_pass2_kernel(const uint2 *__restrict__ lrow4, const float4 *__restrict__ prod4,
                          const int32_t *__restrict__ seg_ptr, int ncb, int nsub, int rs, int n, float *__restrict__ y,
                          float beta, int red) {
    extern __shared__ float y_all[];                        // [warps][rs]
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sub = blockIdx.x * kP2Warps + w;
    if (sub >= nsub) return;                                // warps are independent: no CTA barrier below
    float *y_s = y_all + (size_t)w * rs;
    // sums start at -0: a row without entries then adds nothing ((-0) + p = p, c + (-0) = c, bit for bit)
    const float negz = __int_as_float(0x80000000);
    for (int i = lane; i < rs; i += 32) y_s[i] = negz;
    __syncwarp();
    const size_t stride = (size_t)nsub + 1;
    const int32_t *sp = seg_ptr + sub;                      // (block cb, this sub-block) at sp[cb * stride]
    int pb_next = 0, pe_next = 0;                           // in quads
    if (lane < ncb) { pb_next = sp[lane * stride] >> 2; pe_next = sp[lane * stride + 1] >> 2; }
    for (int cb0 = 0; cb0 < ncb; cb0 += 32) {               // a window of 32 column blocks, one per lane
        const int pb = pb_next, pe = pe_next;
        pb_next = pe_next = 0;
        if (cb0 + 32 + lane < ncb) {                        // the next window's pointers travel while this one is added
            pb_next = sp[(size_t)(cb0 + 32 + lane) * stride] >> 2;
            pe_next = sp[(size_t)(cb0 + 32 + lane) * stride + 1] >> 2;
        }
        const int nc = (pe - pb + 31) >> 5;                 // chunks of this lane's sub-segment
        int cs = nc;                                        // inclusive scan: chunks up to and including this lane
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, cs, o);
            if (lane >= o) cs += t;
        }
        const int total_chunks = __shfl_sync(0xffffffffu, cs, 31);
        // chunk j of the window belongs to the first lane whose inclusive count exceeds j
        auto issue = [&](int j, uint2 &r, float4 &v) {
            r = make_uint2(0xffffffffu, 0xffffffffu);
            v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (j < total_chunks) {                         // uniform
                const int l = __ffs(__ballot_sync(0xffffffffu, cs > j)) - 1;
                const int b = __shfl_sync(0xffffffffu, pb, l), e = __shfl_sync(0xffffffffu, pe, l);
                const int first = __shfl_sync(0xffffffffu, cs - nc, l);
                const int q = b + ((j - first) << 5) + lane;
                if (q < e) { r = __ldcs(lrow4 + q); v = __ldcs(prod4 + q); }
            }
        };
        uint2 rq[D];
        float4 vq[D];
#pragma unroll
        for (int d = 0; d < D; ++d) issue(d, rq[d], vq[d]);
        for (int j0 = 0; j0 < total_chunks; j0 += D) {
#pragma unroll
            for (int d = 0; d < D; ++d) {
                if (j0 + d < total_chunks) {                // uniform
                    const uint2 r = rq[d];
                    const float4 v = vq[d];
                    issue(j0 + d + D, rq[d], vq[d]);
                    const uint32_t r0 = r.x & 0xffffu, r1 = r.x >> 16, r2 = r.y & 0xffffu, r3 = r.y >> 16;
                    const bool e1 = r1 == r0, e2 = r2 == r1, e3 = r3 == r2;
                    const float s0 = v.x;
                    const float s1 = e1 ? s0 + v.y : v.y;
                    const float s2 = e2 ? s1 + v.z : v.z;
                    const float s3 = e3 ? s2 + v.w : v.w;
                    const uint32_t prev3 = __shfl_up_sync(0xffffffffu, r3, 1);
                    const bool cont = lane > 0 && r0 == prev3 && r0 != 0xffffu;      // my head run began in an earlier lane
                    const int h = !e1 ? 0 : (!e2 ? 1 : (!e3 ? 2 : 3));               // slot that closes the head run
                    const unsigned cmask = __ballot_sync(0xffffffffu, cont);
                    // phase 0: every run that closes inside this lane, except a held-back head run
                    if (!e1 && r0 != 0xffffu && !cont) y_s[r0] += s0;
                    if (!e2 && r1 != 0xffffu && !(cont && h == 1)) y_s[r1] += s1;
                    if (!e3 && r2 != 0xffffu && !(cont && h == 2)) y_s[r2] += s2;
                    if (r3 != 0xffffu && !(cont && h == 3)) y_s[r3] += s3;
                    if (cmask) {                            // uniform
                        const float hs = h == 0 ? s0 : (h == 1 ? s1 : (h == 2 ? s2 : s3));
                        // lanes whose whole quad continues a run pass it on: my distance = 1 + such lanes directly below me
                        const unsigned fmask = __ballot_sync(0xffffffffu, cont && h == 3);
                        const unsigned below = lane ? (fmask << (32 - lane)) : 0u;
                        const int dist = cont ? 1 + __clz(~below) : 0;
                        for (int p = 1;; ++p) {
                            __syncwarp();
                            if (dist == p) y_s[r0] += hs;
                            if (!__any_sync(0xffffffffu, dist > p)) break;
                        }
                    }
                    __syncwarp();                           // the next chunk may continue a row of this one
                }
            }
        }
    }
    __syncwarp();
    const int row0 = sub * rs;
    for (int i = lane; i < rs && row0 + i < n; i += 32) {
        const float s = y_s[i];
        const bool nothing = __float_as_uint(s) == 0x80000000u;       // no entry (or only -0 products): the sum changes nothing
        if (red) {                                          // y in a peer GPU's memory, beta == 1: add at the owner, never read
            if (!nothing) red_add_f32(y + row0 + i, s);
            continue;
        }
        if (nothing && beta == 1.0f) continue;              // column without entries, no scale: untouched
        float c0 = y[row0 + i];
        if (beta != 1.0f) c0 *= beta;
        y[row0 + i] = c0 + s;
    }
}

}  // namespace

int run_spmv_blocked(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta, bool red, cudaStream_t stream) {
    const BlockedFormat &bf = mat->blocked;
    static int sms = 0;
    if (!sms && (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mat->device) != cudaSuccess || sms <= 0)) sms = 148;
    const size_t smem1 = (size_t)(256 * 32 + bf.cbw) * 4;
    const size_t smem2 = (size_t)kP2Warps * bf.rs * 4;
    static bool configured_on[64] = {false};
    bool &configured = configured_on[mat->device & 63];
    if (!configured) {
        SB_CUDA(cudaFuncSetAttribute(spmv_blocked_pass1_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        SB_CUDA(cudaFuncSetAttribute(spmv_blocked_pass2_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, kP2Warps * kBlockedMaxRows * 4));
        configured = true;
    }
    struct Timer {                                           // event pair around both passes when profiling
        sblas_b200_matrix *mat; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
        Timer(sblas_b200_matrix *m_, cudaStream_t s_) : mat(m_), st(s_) {
            if (mat->profiling) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, st); }
        }
        ~Timer() { if (a) { cudaEventRecord(b, st); mat->prof_events.push_back(std::make_pair(a, b)); } }
    } timer(mat, stream);
    // 3 B read + 4 B written by pass 1, 6 B read by pass 2 per blocked entry; the long rows stay at 5 B
    note_product(mat, red ? "spmv_blocked red" : "spmv_blocked", 1,
                 (13.0 * (double)bf.entries + 5.0 * (double)(mat->nnz - bf.entries)) / (double)mat->nnz,
                 (int64_t)bf.ncb * (bf.nsub + 1) * 4);
    const int x_vec = (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (bf.cbw & 3) == 0;
    const int grid1 = (int)std::max<int64_t>(1, std::min<int64_t>(sms, bf.total / (4 * kP1Threads * 4)));
    SB_LAUNCH(spmv_blocked_pass1_kernel<4>, grid1, kP1Threads, smem1, stream, reinterpret_cast<const uint2 *>(bf.d_lcol),
              reinterpret_cast<const uint32_t *>(bf.d_code), bf.d_cb_start, bf.ncb, bf.cbw, x, mat->k, mat->d_table, mat->T,
              alpha, reinterpret_cast<float4 *>(bf.d_prod), x_vec);
    const int grid2 = (bf.nsub + kP2Warps - 1) / kP2Warps;
    SB_LAUNCH(spmv_blocked_pass2_kernel<4>, grid2, 32 * kP2Warps, smem2, stream, reinterpret_cast<const uint2 *>(bf.d_lrow),
              reinterpret_cast<const float4 *>(bf.d_prod), bf.d_seg_ptr, bf.ncb,
              bf.nsub, bf.rs, mat->n, y, beta, red ? 1 : 0);
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200
