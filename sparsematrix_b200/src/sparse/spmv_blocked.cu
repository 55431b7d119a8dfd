// spmv_blocked.cu -- K-spmv-blocked: SpMV (m == 1) in two streaming passes over the column-blocked
// copy of W (src/utils/blocked_format.cu), for matrices whose x gathers would each cost an L2 sector.
//
// Replaces, like the row kernels of product.cu, the reference's decode + AXPY pair at m = 1
// (kernel.cc:771-800 with the scalar tail :568-582): y[j] = beta*y[j] + sum_r x[r] * fl(table[code]*alpha).
//
//   pass 1  persistent CTAs walk the slot stream; the x tile of the current column block (cbw floats)
//           sits in shared memory, the codebook * alpha is replicated per lane (conflict-free look-up):
//           p[slot] = fl(x_s[lcol] * tab[code]), and inside a quad of slots the products of one row's run are
//           added left to right (the format keeps a run inside a quad): 3 B read + 4 B written per slot,
//           no global gather
//   pass 2  one warp per row sub-block (rs rows, sums in shared memory) walks its sub-segments in column
//           block order, four neighbouring slots per lane (8 + 16 byte loads), and adds every slot that closes
//           a run to its row's sum: 6 B read per slot.  Within one 128-slot step no two lanes hold the same row
//           (a row closes one run per segment) -- except for the quads of a run longer than a quad, which carry a
//           flag and are added one lane after the other.  No CTA barrier anywhere: the warps of a CTA own disjoint rows.
// Products and sums are separately rounded like the reference's generic build (no FMA); a row's terms are added
// run by run in column order, the terms of a run first among themselves: fixed order, run-to-run identical.
// HBM traffic: 13 B per slot + x + y, against 5 B per entry and one L2 sector per entry for the row kernels.
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
                          const int32_t *__restrict__ cb_start, int cb_lo, int cb_hi, int cbw, const float *__restrict__ x, int k,
                          const float *__restrict__ table, int T, float alpha, float4 *__restrict__ prod4, int x_vec) {
    extern __shared__ __align__(16) float smem[];
    float *tab_s = smem;                                    // [256][32]: entry c of lane l at c * 32 + l
    float *x_s = smem + 256 * 32;                           // [cbw]
    const int lane = threadIdx.x & 31;
    {
        // one global load per thread, eight of the 32 copies of its entry each (a loop over the 8192 words was eight
        // dependent loads per thread at the head of the kernel)
        static_assert(kP1Threads == 1024, "four threads per codebook entry");
        const int c = threadIdx.x >> 2;
        const float v = c < T ? __ldg(table + c) * alpha : 0.0f;    // val_table[code] * alpha (kernel.cc:791)
        const float4 q = make_float4(v, v, v, v);
        float4 *dst = reinterpret_cast<float4 *>(tab_s + c * 32 + (threadIdx.x & 3) * 8);
        dst[0] = q;
        dst[1] = q;
    }
    // this launch covers the column blocks [cb_lo, cb_hi): quads [q_lo, q_hi), a contiguous share per CTA
    const int q_lo = cb_start[cb_lo] >> 2, total_q = (cb_start[cb_hi] >> 2) - q_lo;
    constexpr int kStep = kP1Threads * U;
    const int q_first = q_lo + (int)(((int64_t)total_q * blockIdx.x / gridDim.x) / kStep * kStep);
    const int q_end = q_lo + (blockIdx.x + 1 == gridDim.x ? total_q
                                                          : (int)(((int64_t)total_q * (blockIdx.x + 1) / gridDim.x) / kStep * kStep));
    int q = q_first, cb = cb_lo;
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
                    float4 p;                               // bit 15 of a local column: the slot continues its left neighbour's run
                    p.x = x_s[lc[u].x & 0x7fffu] * tab_l[(cd[u] << 5) & 0x1fe0u];
                    p.y = x_s[(lc[u].x >> 16) & 0x7fffu] * tab_l[(cd[u] >> 3) & 0x1fe0u];
                    p.z = x_s[lc[u].y & 0x7fffu] * tab_l[(cd[u] >> 11) & 0x1fe0u];
                    p.w = x_s[(lc[u].y >> 16) & 0x7fffu] * tab_l[(cd[u] >> 19) & 0x1fe0u];
                    if (lc[u].x & 0x80000000u) p.y = p.x + p.y;
                    if (lc[u].y & 0x00008000u) p.z = p.y + p.z;
                    if (lc[u].y & 0x80000000u) p.w = p.z + p.w;
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

// One step of a warp is a chunk of 32 quads (128 slots) of one sub-segment: lane l holds four neighbouring slots -- their
// row fields (0xffff: nothing to add; bit 14: closes a run that began in an earlier quad) and their pass-1 sums.  Slots
// without bit 14 name pairwise different rows across the whole chunk, so the lanes add them at once, slot index after slot
// index; the flagged ones (at most one per lane) follow, one lane after the other in lane = column order.
// D = chunks a warp keeps in flight.
template <int D>
__global__ void __launch_bounds__(32 * kP2Warps, 2)
spmv_blocked_pass2_kernel(const uint2 *__restrict__ lrow4, const float4 *__restrict__ prod4,
                          const int32_t *__restrict__ seg_ptr, int cb_lo, int ncb, int nsub, int rs, int n, float *__restrict__ y,
                          float beta, int red, int abl) {      // column blocks [cb_lo, ncb)
    extern __shared__ float y_all[];                        // [warps][rs]
    float sink = 0.0f;                                      // experiments (abl != 0): where the sums go instead of shared memory
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
    if (cb_lo + lane < ncb) { pb_next = sp[(size_t)(cb_lo + lane) * stride] >> 2; pe_next = sp[(size_t)(cb_lo + lane) * stride + 1] >> 2; }
    // One chunk into the sums.  The fields without bit 14 name pairwise different rows across the chunk (also inside a
    // lane): four independent read-modify-writes per lane, reads first.
    auto add_chunk = [&](const uint2 r, const float4 v) {
        if (abl) { sink += v.x + v.y + v.z + v.w + __uint_as_float(r.x ^ r.y); return; }   // timing experiment
        const uint32_t late = ((r.x & ~(r.x >> 1)) | (r.y & ~(r.y >> 1))) & 0x40004000u;
        const unsigned late_lanes = __ballot_sync(0xffffffffu, late != 0u);
        const bool a0 = !(r.x & 0x0000c000u), a1 = !(r.x & 0xc0000000u), a2 = !(r.y & 0x0000c000u), a3 = !(r.y & 0xc0000000u);
        float *p0 = y_s + (r.x & 0x3fffu), *p1 = y_s + ((r.x >> 16) & 0x3fffu), *p2 = y_s + (r.y & 0x3fffu),
              *p3 = y_s + ((r.y >> 16) & 0x3fffu);
        float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f;
        if (a0) t0 = *p0;
        if (a1) t1 = *p1;
        if (a2) t2 = *p2;
        if (a3) t3 = *p3;
        t0 += v.x; t1 += v.y; t2 += v.z; t3 += v.w;
        if (a0) *p0 = t0;
        if (a1) *p1 = t1;
        if (a2) *p2 = t2;
        if (a3) *p3 = t3;
        if (late_lanes) {                                   // uniform; rare: a row with more than four entries in one column block
            uint32_t lr = 0xffffu;                          // the lane's flagged field, if any (at most one per quad)
            float lv = 0.0f;
            if ((r.x & 0x0000c000u) == 0x00004000u) { lr = r.x & 0xffffu; lv = v.x; }
            if ((r.x & 0xc0000000u) == 0x40000000u) { lr = r.x >> 16; lv = v.y; }
            if ((r.y & 0x0000c000u) == 0x00004000u) { lr = r.y & 0xffffu; lv = v.z; }
            if ((r.y & 0xc0000000u) == 0x40000000u) { lr = r.y >> 16; lv = v.w; }
            for (unsigned todo = late_lanes; todo; todo &= todo - 1) {
                __syncwarp();
                if (lane == __ffs(todo) - 1) y_s[lr & 0x3fffu] += lv;
            }
        }
        __syncwarp();                                       // the next chunk may name a row of this one
    };
    for (int cb0 = cb_lo; cb0 < ncb; cb0 += 32) {           // a window of 32 column blocks, one per lane
        const int pb = pb_next, pe = pe_next;
        pb_next = pe_next = 0;
        if (cb0 + 32 + lane < ncb) {                        // the next window's pointers travel while this one is added
            pb_next = sp[(size_t)(cb0 + 32 + lane) * stride] >> 2;
            pe_next = sp[(size_t)(cb0 + 32 + lane) * stride + 1] >> 2;
        }
        const int ncw = min(32, ncb - cb0);
        // the first 32 quads of column block cb0 + c (nearly always the whole sub-segment)
        auto fetch = [&](int c, uint2 &r, float4 &v) {
            r = make_uint2(0xffffffffu, 0xffffffffu);
            v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c < ncw) {                                  // uniform
                const int q = __shfl_sync(0xffffffffu, pb, c) + lane;
                if (q < __shfl_sync(0xffffffffu, pe, c)) { r = __ldcs(lrow4 + q); v = __ldcs(prod4 + q); }
            }
        };
        // D sub-segments per step: their loads go out together, then they are added one after the other.  (A load issued
        // while older ones are awaited buys nothing here: the wait is for every load of the warp that is in flight.)
        for (int c0 = 0; c0 < ncw; c0 += D) {
            uint2 rq[D];
            float4 vq[D];
#pragma unroll
            for (int d = 0; d < D; ++d) fetch(c0 + d, rq[d], vq[d]);
#pragma unroll
            for (int d = 0; d < D; ++d) {
                const int c = c0 + d;
                if (c < ncw) {                              // uniform
                    add_chunk(rq[d], vq[d]);
                    const int e = __shfl_sync(0xffffffffu, pe, c);
                    for (int q0 = __shfl_sync(0xffffffffu, pb, c) + 32; q0 < e; q0 += 32) {   // a sub-segment of more than 128 slots
                        uint2 r2 = make_uint2(0xffffffffu, 0xffffffffu);
                        float4 v2 = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (q0 + lane < e) { r2 = __ldcs(lrow4 + q0 + lane); v2 = __ldcs(prod4 + q0 + lane); }
                        add_chunk(r2, v2);
                    }
                }
            }
        }
    }
    __syncwarp();
    if (abl && sink == 12345.678f) y_s[0] = sink;
    const int row0 = sub * rs;
    const int rows = min(rs, n - row0);
    constexpr int EU = 8;                                   // rows per lane and step: their y values travel together
    for (int i0 = lane; i0 < rows; i0 += 32 * EU) {
        float s[EU], c0[EU];
#pragma unroll
        for (int u = 0; u < EU; ++u) {
            const int i = i0 + 32 * u;
            s[u] = i < rows ? y_s[i] : negz;
            // no entry (or only -0 products): the sum changes nothing; y is read only where it will be rewritten
            const bool nothing = __float_as_uint(s[u]) == 0x80000000u;
            c0[u] = 0.0f;
            if (i < rows && !red && !(nothing && beta == 1.0f)) c0[u] = y[row0 + i];
        }
#pragma unroll
        for (int u = 0; u < EU; ++u) {
            const int i = i0 + 32 * u;
            if (i >= rows) continue;
            const bool nothing = __float_as_uint(s[u]) == 0x80000000u;
            if (red) {                                      // y in a peer GPU's memory, beta == 1: add at the owner, never read
                if (!nothing) red_add_f32(y + row0 + i, s[u]);
                continue;
            }
            if (nothing && beta == 1.0f) continue;          // column without entries, no scale: untouched
            float c = c0[u];
            if (beta != 1.0f) c *= beta;
            y[row0 + i] = c + s[u];
        }
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
        SB_CUDA(cudaFuncSetAttribute(spmv_blocked_pass2_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kP2Warps * kBlockedMaxRows * 4));
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
    // experiment switches (timing only; results are wrong with them): SBLAS_B200_BLK_ONLY = 1 | 2 runs one pass,
    // SBLAS_B200_BLK_ABL = 1 drops pass 2's shared-memory updates; SBLAS_B200_BLK_D = 2 | 4 sub-segments per pass-2 step
    // (profiles/r02_blocked_spmv.md)
    const char *e_only = getenv("SBLAS_B200_BLK_ONLY"), *e_abl = getenv("SBLAS_B200_BLK_ABL");
    const int only = e_only ? atoi(e_only) : 0, abl = e_abl ? atoi(e_abl) : 0;
    const int grid2 = (bf.nsub + kP2Warps - 1) / kP2Warps;
    const char *e_d = getenv("SBLAS_B200_BLK_D");
    const int depth = e_d ? atoi(e_d) : 2;
    const int grid1 = (int)std::max<int64_t>(1, std::min<int64_t>(sms, bf.total / (4 * kP1Threads * 4)));
    if (only != 2)
        SB_LAUNCH(spmv_blocked_pass1_kernel<4>, grid1, kP1Threads, smem1, stream, reinterpret_cast<const uint2 *>(bf.d_lcol),
                  reinterpret_cast<const uint32_t *>(bf.d_code), bf.d_cb_start, 0, bf.ncb, bf.cbw, x, mat->k, mat->d_table, mat->T,
                  alpha, reinterpret_cast<float4 *>(bf.d_prod), x_vec);
    // (pass 2 launched programmatically dependent on pass 1 was measured: the cold step went from 410 to 446 us, r02)
#define SB_P2(DD) SB_LAUNCH(spmv_blocked_pass2_kernel<DD>, grid2, 32 * kP2Warps, smem2, stream, reinterpret_cast<const uint2 *>(bf.d_lrow), \
              reinterpret_cast<const float4 *>(bf.d_prod), bf.d_seg_ptr, 0, bf.ncb, bf.nsub, bf.rs, mat->n, y, beta, red ? 1 : 0, abl)
    if (only != 1) {
        if (depth == 4) SB_P2(4); else SB_P2(2);
    }
#undef SB_P2
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200
