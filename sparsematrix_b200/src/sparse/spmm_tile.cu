// spmm_tile.cu -- K-spmm (tile form): SpMM with the dense operand staged through shared memory.
//
// Y[n x m] = W X.  A CTA owns nwarps*8 W rows and a tile of MT = 32*NF4 dense columns; it walks
// the X rows in chunks of 255.  Chunk c of the X tile (255 x MT fp32, contiguous in the tile-major
// X^T workspace) is brought in by ONE bulk asynchronous copy (cp.async.bulk -> mbarrier complete_tx,
// the TMA engine; no thread moves data) issued by a producer warp into a ring of STAGES buffers.
// Consumer warps never meet at a CTA barrier inside the k loop: a stage is handed back through an
// "empty" mbarrier (one arrival per warp), so a warp with short rows runs ahead by up to STAGES-1
// chunks instead of idling at every chunk.
//
// A warp owns one group of 8 W rows (src/utils/tile_format.cu) and is arranged as 4 slots x 8
// column lanes: slot s walks rows 8g+s and 8g+4+s, one after the other, each in column order -- the
// order in which the reference adds the terms of one output element (kernel.cc:780-796); column
// lane c holds float4 #c (and #c+8 when NF4 = 2) of the MT dense columns.  One step of a warp is
// one entry per slot: per lane one LDS.128 per float4 and 4 FFMA.  Slots whose row has run out of
// entries in this chunk are predicated off, so they cost neither shared-memory bandwidth nor FMAs.
// This replaces the reference's AXPY micro-kernel (kernel.cc:371-584: cc[0..m) += aa[0..m) * val)
// with the roles of "contiguous m" kept: the m dense values of one X row are contiguous in shared
// memory exactly like aa[] is contiguous there.
//
// The warp's entries for all chunks form one contiguous stream of 80-byte quad rows, which also
// arrives by bulk copies, into a private two-half ring in shared memory (a warp's entry fetch would
// otherwise sit on the L2/HBM latency).
//
// Accumulators stay in registers across the whole k loop (2 rows x NF4 float4 per lane); the result
// tile goes through shared memory so that C (row-major m x n = Y^T) is written in 128-byte rows --
// the function of the reference's per-panel C transposes (sparse-matrix.cc:184,189).  beta is
// applied there (only when != 1, sparse-matrix.cc:149), alpha is folded into the weights first
// (kernel.cc:791).
#include <stdlib.h>

#include <algorithm>

#include "device_matrix.h"

namespace sblas_b200 {

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy, completion counted in bytes on an mbarrier (TMA engine; SASS UBLKCP)
__device__ __forceinline__ void bulk_g2s_addr(uint32_t dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    bulk_g2s_addr(smem_u32(dst), src, bytes, bar);
}

constexpr int kRingHalf = 32;                                   // quad rows per half of a warp's ring
constexpr int kRingBytes = 2 * kRingHalf * kBlockBytes;        // 5120 B per warp
constexpr int kMaxWarps = 16;                                   // consumer warps per CTA

// The warp has just read the last quad row of a ring half (`consumed` read so far): refill that
// half with the quad rows one ring ahead, then make sure the half it moves into has landed.
__device__ __noinline__ void ring_turn(uint32_t ring, uint64_t *wbar, const uint8_t *stream, int consumed, int total,
                                       int lane) {
    const int done_half = ((consumed - 1) / kRingHalf) & 1;
    __syncwarp();
    const int nxt = consumed + kRingHalf;                        // first quad row of the refill
    if (lane == 0 && nxt < total) {
        const int cnt = min(kRingHalf, total - nxt);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&wbar[done_half], (uint32_t)cnt * kBlockBytes);
        bulk_g2s_addr(ring + done_half * kRingHalf * kBlockBytes, stream + (size_t)nxt * kBlockBytes,
                      (uint32_t)cnt * kBlockBytes, &wbar[done_half]);
    }
    if (consumed < total) mbar_wait(&wbar[done_half ^ 1], (uint32_t)((consumed / (2 * kRingHalf)) & 1));
}

// One entry of this lane's slot: acc[0..NF4) += w * X[off][float4 lanes], skipped when off == 255.
// Written as predicated PTX so that idle slots issue no shared-memory access at all.  The staging
// registers x[] are the caller's (one set per step of a quad row, reused from quad row to quad
// row): a predicated load keeps the old value, so private temporaries would stay live for ever.
template <int NF4, int ROW_BYTES>
__device__ __forceinline__ void slot_step(float4 (&acc)[NF4], float4 (&x)[NF4], float w, uint32_t off, uint32_t xbase) {
    if (NF4 == 1) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            ".reg .u32 a;\n"
            "setp.ne.u32 p, %8, 255;\n"
            "mad.lo.u32 a, %8, %9, %10;\n"
            "@p ld.shared.v4.f32 {%4, %5, %6, %7}, [a];\n"
            "@p fma.rn.f32 %0, %11, %4, %0;\n"
            "@p fma.rn.f32 %1, %11, %5, %1;\n"
            "@p fma.rn.f32 %2, %11, %6, %2;\n"
            "@p fma.rn.f32 %3, %11, %7, %3;\n"
            "}\n"
            : "+f"(acc[0].x), "+f"(acc[0].y), "+f"(acc[0].z), "+f"(acc[0].w), "+f"(x[0].x), "+f"(x[0].y), "+f"(x[0].z),
              "+f"(x[0].w)
            : "r"(off), "n"(ROW_BYTES), "r"(xbase), "f"(w));
    } else {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            ".reg .u32 a;\n"
            "setp.ne.u32 p, %16, 255;\n"
            "mad.lo.u32 a, %16, %17, %18;\n"
            "@p ld.shared.v4.f32 {%8, %9, %10, %11}, [a];\n"
            "@p ld.shared.v4.f32 {%12, %13, %14, %15}, [a+128];\n"
            "@p fma.rn.f32 %0, %19, %8, %0;\n"
            "@p fma.rn.f32 %1, %19, %9, %1;\n"
            "@p fma.rn.f32 %2, %19, %10, %2;\n"
            "@p fma.rn.f32 %3, %19, %11, %3;\n"
            "@p fma.rn.f32 %4, %19, %12, %4;\n"
            "@p fma.rn.f32 %5, %19, %13, %5;\n"
            "@p fma.rn.f32 %6, %19, %14, %6;\n"
            "@p fma.rn.f32 %7, %19, %15, %7;\n"
            "}\n"
            : "+f"(acc[0].x), "+f"(acc[0].y), "+f"(acc[0].z), "+f"(acc[0].w), "+f"(acc[NF4 - 1].x),
              "+f"(acc[NF4 - 1].y), "+f"(acc[NF4 - 1].z), "+f"(acc[NF4 - 1].w), "+f"(x[0].x), "+f"(x[0].y),
              "+f"(x[0].z), "+f"(x[0].w), "+f"(x[NF4 - 1].x), "+f"(x[NF4 - 1].y), "+f"(x[NF4 - 1].z),
              "+f"(x[NF4 - 1].w)
            : "r"(off), "n"(ROW_BYTES), "r"(xbase), "f"(w));
    }
}

}  // namespace

// ---- K-transpose-in (tile-major): xt[tile][r][il] = A[tile*MT + il][r], zero beyond m ----------
__global__ void __launch_bounds__(256)
transpose_in_tiled_kernel(const float *__restrict__ a, int m, int k, int lda, float *__restrict__ xt, int k_alloc,
                          int MT) {
    __shared__ float tile[32][33];
    const int r0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int ii = ty; ii < 32; ii += 8) {
        const int i = i0 + ii, r = r0 + tx;
        tile[ii][tx] = (i < m && r < k) ? a[(size_t)i * lda + r] : 0.0f;
    }
    __syncthreads();
    const int i = i0 + tx;                              // dense column; MT is a multiple of 32
    const int t = i / MT, il = i - t * MT;
    for (int rr = ty; rr < 32; rr += 8) {
        const int r = r0 + rr;
        if (r < k) xt[((size_t)t * k_alloc + r) * MT + il] = tile[tx][rr];
    }
}

// blockDim.x = (nwarps + 1) * 32: nwarps consumer warps and one producer warp (the last).
template <int NF4, int STAGES, bool SCALE>
__global__ void __maxnreg__(112)
spmm_tile_kernel(const uint8_t *__restrict__ blocks, const int32_t *__restrict__ grp_ptr,
                 const uint32_t *__restrict__ grp_nq, const int32_t *__restrict__ row_ptr, int ngroups, int nchunks,
                 int n, const float *__restrict__ xt, int k_alloc, float *__restrict__ c, int ldc, int m, float alpha,
                 float beta, int nwarps) {
    constexpr int MT = 32 * NF4;
    constexpr int STAGE_BYTES = 256 * MT * 4;                    // 255 data rows, padded to a round size
    constexpr uint32_t CHUNK_BYTES = kTileK * MT * 4;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t *ring_base = smem_raw + (size_t)STAGES * STAGE_BYTES;                           // [nwarps][kRingBytes]
    uint64_t *full = reinterpret_cast<uint64_t *>(ring_base + (size_t)nwarps * kRingBytes);  // [STAGES]
    uint64_t *empty = full + STAGES;                                                         // [STAGES]
    uint64_t *wfull = empty + STAGES;                                                        // [nwarps][2]
    __shared__ int s_len[kMaxWarps * kTileRows];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = lane >> 3, c8 = lane & 7;
    const int RB = nwarps * kTileRows;
    const int j0 = blockIdx.x * RB;
    const float *xt_tile = xt + (size_t)blockIdx.y * k_alloc * MT;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], nwarps); }
        for (int s = 0; s < 2 * nwarps; ++s) mbar_init(&wfull[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < RB; i += blockDim.x) {
        const int j = j0 + i;
        s_len[i] = j < n ? row_ptr[j + 1] - row_ptr[j] : 0;
    }
    __syncthreads();

    float4 acc[2][NF4];
#pragma unroll
    for (int rr = 0; rr < 2; ++rr)
#pragma unroll
        for (int f = 0; f < NF4; ++f) acc[rr][f] = make_float4(0.f, 0.f, 0.f, 0.f);

    if (warp == nwarps) {
        // ---- producer: keep the X ring full ----
        if (lane == 0) {
            for (int ch = 0; ch < nchunks; ++ch) {
                const int stage = ch % STAGES, round = ch / STAGES;
                if (round > 0) mbar_wait(&empty[stage], (uint32_t)((round - 1) & 1));
                mbar_expect_tx(&full[stage], CHUNK_BYTES);
                bulk_g2s(smem_raw + (size_t)stage * STAGE_BYTES, xt_tile + (size_t)ch * kTileK * MT, CHUNK_BYTES,
                         &full[stage]);
            }
        }
    } else {
        // ---- consumers ----
        const int g = blockIdx.x * nwarps + warp;
        const bool live = g < ngroups;
        // The warp's quad-row stream -- contiguous over all chunks of its group -- flows through a
        // private two-half ring.  Each half is filled by one bulk copy; when the warp has read the
        // last quad row of a half, lane 0 refills that half (ring_turn, off the hot path).
        const int64_t s0 = live ? grp_ptr[(int64_t)g * nchunks] : 0;
        const int total = live ? (int)(grp_ptr[(int64_t)(g + 1) * nchunks] - s0) : 0;
        const uint8_t *ring_ptr = ring_base + (size_t)warp * kRingBytes;
        const uint32_t ring = smem_u32(ring_ptr);
        uint64_t *wbar = wfull + 2 * warp;
        const uint8_t *stream = blocks + s0 * kBlockBytes;
        if (lane == 0) {
            for (int h = 0; h < 2; ++h) {
                const int cnt = min(kRingHalf, total - h * kRingHalf);
                if (cnt > 0) {
                    mbar_expect_tx(&wbar[h], (uint32_t)cnt * kBlockBytes);
                    bulk_g2s_addr(ring + h * kRingHalf * kBlockBytes, stream + (size_t)h * kRingHalf * kBlockBytes,
                                  (uint32_t)cnt * kBlockBytes, &wbar[h]);
                }
            }
        }
        // One quad row is always held in registers ahead of the one being used; `reads` counts the
        // quad rows taken out of the ring, `left` those still unread in the half `rp` points into.
        // Reading one quad row past the end of the stream is harmless (stale ring bytes, never used).
        int reads = 0, left = kRingHalf;
        const uint8_t *rp = ring_ptr + slot * 16;                // this slot's read pointer into the ring
        const int roff = 64 - slot * 12;                         // from the slot's weights to its offsets
        if (total > 0) mbar_wait(&wbar[0], 0u);
        float4 cur_w = *reinterpret_cast<const float4 *>(rp);
        uint32_t cur_r = *reinterpret_cast<const uint32_t *>(rp + roff);
        rp += kBlockBytes;
        reads = 1;
        left = kRingHalf - 1;
        uint32_t nq_next = live ? grp_nq[(int64_t)g * nchunks] : 0u;
        const uint32_t xlane = smem_u32(smem_raw) + c8 * 16;

        float4 xr[4][NF4];                                       // X staging registers, one set per step
#pragma unroll
        for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int f = 0; f < NF4; ++f) xr[t][f] = make_float4(0.f, 0.f, 0.f, 0.f);
        // the quad rows of one row pair in the current chunk
        auto walk_pair = [&](float4 (&a)[NF4], int nq, uint32_t xbase) {
            while (nq > 0) {                                     // runs that stay inside one ring half
                const int run = min(nq, left);
#pragma unroll 1
                for (int q = 0; q < run; ++q) {
                    float4 w = cur_w;
                    const uint32_t r4 = cur_r;
                    cur_w = *reinterpret_cast<const float4 *>(rp);              // next quad row, ahead of use
                    cur_r = *reinterpret_cast<const uint32_t *>(rp + roff);
                    rp += kBlockBytes;
                    if (SCALE) { w.x *= alpha; w.y *= alpha; w.z *= alpha; w.w *= alpha; }
                    slot_step<NF4, MT * 4>(a, xr[0], w.x, __byte_perm(r4, 0u, 0x4440u), xbase);
                    slot_step<NF4, MT * 4>(a, xr[1], w.y, __byte_perm(r4, 0u, 0x4441u), xbase);
                    slot_step<NF4, MT * 4>(a, xr[2], w.z, __byte_perm(r4, 0u, 0x4442u), xbase);
                    slot_step<NF4, MT * 4>(a, xr[3], w.w, __byte_perm(r4, 0u, 0x4443u), xbase);
                }
                nq -= run;
                left -= run;
                reads += run;
                if (left == 0) {                                 // the look-ahead read was the last of a half
                    ring_turn(ring, wbar, stream, reads, total, lane);
                    left = kRingHalf;
                    if (((reads / kRingHalf) & 1) == 0) rp = ring_ptr + slot * 16;
                }
            }
        };

        for (int ch = 0; ch < nchunks; ++ch) {
            const int stage = ch % STAGES;
            const uint32_t nq2 = nq_next;
            if (live && ch + 1 < nchunks) nq_next = grp_nq[(int64_t)g * nchunks + ch + 1];
            mbar_wait(&full[stage], (uint32_t)((ch / STAGES) & 1));
            const uint32_t xbase = xlane + stage * STAGE_BYTES;
            walk_pair(acc[0], (int)(nq2 & 0xffffu), xbase);
            walk_pair(acc[1], (int)(nq2 >> 16), xbase);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);           // this warp is done with the stage
        }
    }

    // every warp is past the k loop: the X ring is free, stage the RB x MT result tile in it
    __syncthreads();
    float *tile = reinterpret_cast<float *>(smem_raw);           // [RB][MT + 1]
    if (warp < nwarps) {
#pragma unroll
        for (int rr = 0; rr < 2; ++rr)
#pragma unroll
            for (int f = 0; f < NF4; ++f) {
                float *p = tile + (size_t)(warp * kTileRows + rr * 4 + slot) * (MT + 1) + 32 * f + c8 * 4;
                p[0] = acc[rr][f].x; p[1] = acc[rr][f].y; p[2] = acc[rr][f].z; p[3] = acc[rr][f].w;
            }
    }
    __syncthreads();
    const int m0 = blockIdx.y * MT;
    const int nw_all = blockDim.x >> 5;
    for (int i = warp; i < MT; i += nw_all) {
        const int gi = m0 + i;
        if (gi >= m) break;
        for (int jl = lane; jl < RB; jl += 32) {
            const int j = j0 + jl;
            if (j >= n) break;
            const int len = s_len[jl];
            if (len == 0 && beta == 1.0f) continue;              // S column without entries, no scale: untouched
            float *p = c + (size_t)gi * ldc + j;
            float c0 = *p;
            if (beta != 1.0f) c0 *= beta;
            *p = len ? c0 + tile[(size_t)jl * (MT + 1) + i] : c0;
        }
    }
}

template <int NF4, int STAGES>
static int launch_tile(sblas_b200_matrix *mat, const float *xt, int k_alloc, float *d_c, int ldc, int m, float alpha,
                       float beta, int nwarps, cudaStream_t stream) {
    constexpr int MT = 32 * NF4;
    const size_t smem = (size_t)STAGES * 256 * MT * 4 + (size_t)nwarps * kRingBytes + (2 * STAGES + 2 * nwarps) * 8;
    static bool configured = false;                              // per template instance
    if (!configured) {
        const int most = (int)((size_t)STAGES * 256 * MT * 4 + (size_t)kMaxWarps * kRingBytes + (2 * STAGES + 2 * kMaxWarps) * 8);
        SB_CUDA(cudaFuncSetAttribute(spmm_tile_kernel<NF4, STAGES, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, most));
        SB_CUDA(cudaFuncSetAttribute(spmm_tile_kernel<NF4, STAGES, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, most));
        configured = true;
    }
    const dim3 grid((mat->tile_ngroups + nwarps - 1) / nwarps, (m + MT - 1) / MT);
    if (alpha == 1.0f)
        SB_LAUNCH((spmm_tile_kernel<NF4, STAGES, false>), grid, (nwarps + 1) * 32, smem, stream, mat->d_tile_blocks,
                  mat->d_grp_ptr, mat->d_grp_nq, mat->d_row_ptr, mat->tile_ngroups, mat->tile_nchunks, mat->n, xt,
                  k_alloc, d_c, ldc, m, alpha, beta, nwarps);
    else
        SB_LAUNCH((spmm_tile_kernel<NF4, STAGES, true>), grid, (nwarps + 1) * 32, smem, stream, mat->d_tile_blocks,
                  mat->d_grp_ptr, mat->d_grp_nq, mat->d_row_ptr, mat->tile_ngroups, mat->tile_nchunks, mat->n, xt,
                  k_alloc, d_c, ldc, m, alpha, beta, nwarps);
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

// Row block height (consumer warps per CTA) that fills the SMs best: one CTA per SM at a time, so
// the cost of a launch is (waves of CTAs) x (rows per CTA); fewer, taller CTAs re-read X less often.
static int pick_nwarps(int ngroups, int mtiles, int sms) {
    int best = kMaxWarps;
    double best_cost = 1e30;
    for (int nw = kMaxWarps; nw >= 4; --nw) {
        const int64_t ctas = (int64_t)((ngroups + nw - 1) / nw) * mtiles;
        const int64_t waves = (ctas + sms - 1) / sms;
        const double cost = (double)waves * (nw + 0.75);        // +0.75: per-CTA X fill and prologue/epilogue
        if (cost < best_cost - 1e-9) { best_cost = cost; best = nw; }
    }
    return best;
}

// Variant selection: SBLAS_B200_TILE = "<warps: 4..g(16), hex><float4s per lane: 1|2>" (experiments), else by shape.
int run_spmm_tile(sblas_b200_matrix *mat, const float *d_a, int m, int lda, float *d_c, int ldc, float alpha,
                  float beta, cudaStream_t stream) {
    static int sms = 0;
    if (!sms) {
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mat->device) != cudaSuccess || sms <= 0) sms = 148;
    }
    // two float4 per lane halve the per-entry bookkeeping, but only when that still fills the SMs
    int nf4 = 1;
    if (m > 32 && (int64_t)((mat->tile_ngroups + 7) / 8) * ((m + 63) / 64) >= sms) nf4 = 2;
    int nwarps = 0;
    if (const char *env = getenv("SBLAS_B200_TILE")) {
        if (env[0] >= '4' && env[0] <= '9') nwarps = env[0] - '0';
        else if (env[0] >= 'a' && env[0] <= 'g') nwarps = 10 + (env[0] - 'a');
        if (env[0] && (env[1] == '1' || env[1] == '2')) nf4 = env[1] - '0';
    }
    const int MT = 32 * nf4;
    const int ntiles = (m + MT - 1) / MT;
    if (nwarps == 0) nwarps = pick_nwarps(mat->tile_ngroups, ntiles, sms);
    const int k_alloc = mat->tile_nchunks * kTileK;
    int rc = ensure_buffer(&mat->d_ws, &mat->ws_bytes, (size_t)ntiles * k_alloc * MT * 4);
    if (rc != SBLAS_B200_OK) return rc;
    const dim3 tgrid((mat->k + 31) / 32, ntiles * MT / 32);
    SB_LAUNCH(transpose_in_tiled_kernel, tgrid, 256, 0, stream, d_a, m, mat->k, lda, mat->d_ws, k_alloc, MT);
    SB_CHECK_LAUNCH();
    struct Timer {                                                // event pair around the dominant kernel when profiling
        sblas_b200_matrix *mat; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
        Timer(sblas_b200_matrix *m_, cudaStream_t s_) : mat(m_), st(s_) {
            if (mat->profiling) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, st); }
        }
        ~Timer() { if (a) { cudaEventRecord(b, st); mat->prof_events.push_back(std::make_pair(a, b)); } }
    } timer(mat, stream);
    if (nf4 == 2) return launch_tile<2, 2>(mat, mat->d_ws, k_alloc, d_c, ldc, m, alpha, beta, nwarps, stream);
    return launch_tile<1, 3>(mat, mat->d_ws, k_alloc, d_c, ldc, m, alpha, beta, nwarps, stream);
}

}  // namespace sblas_b200
