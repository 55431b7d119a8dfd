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
// The warp's entries for all chunks form one contiguous stream of 80-byte quad rows, which the
// warp copies into a private ring in shared memory with cp.async, three segments of 8 quad rows
// ahead of use (a warp's entry fetch would otherwise sit on the L2/HBM latency).
//
// Accumulators stay in registers across the whole k loop (2 rows x NF4 float4 per lane); the result
// tile goes through shared memory so that C (row-major m x n = Y^T) is written in 128-byte rows --
// the function of the reference's per-panel C transposes (sparse-matrix.cc:184,189).  beta is
// applied there (only when != 1, sparse-matrix.cc:149), alpha is folded into the weights first
// (kernel.cc:791).
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "async_copy.cuh"
#include "device_matrix.h"

namespace sblas_b200 {

namespace {

#ifndef TILE_MAX_ROWS
#define TILE_MAX_ROWS 128
#endif
#ifndef TILE_MAXNREG_BIG
#define TILE_MAXNREG_BIG 112
#endif
#ifndef TILE_MAXW_BIG
#define TILE_MAXW_BIG 16
#endif
constexpr int kMaxRows = TILE_MAX_ROWS;                                   // W rows per CTA at most

#ifdef EXP_TIMELINE
__device__ unsigned long long g_timeline[2048 * 8];
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define TL(p) do { if (threadIdx.x == 0) g_timeline[blockIdx.x * 8 + (p)] = gtime(); } while (0)
#else
#define TL(p) do { } while (0)
#endif

// The warp's quad-row stream goes through a private ring of kRingSegs segments of kSegRows quad rows;
// a segment is fetched by the warp itself with 16-byte cp.async (LDGSTS) three segments ahead.
constexpr int kSegRows = 8, kRingSegs = 4;
// epilogue variants (kernel argument `epi`): C in local memory is prefetched into L2 during the k loop and
// read-modify-written; C in a peer GPU's memory is not prefetched and, for beta == 1, updated by reductions
constexpr int kEpiNoPrefetch = 1, kEpiRedAdd = 2;
constexpr int kSegBytes = kSegRows * kBlockBytes;                // 640
constexpr int kRingBytes = kRingSegs * kSegBytes;                // 2560 B per warp

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// fetch segment `seg` of the stream into its ring position; always one commit group
__device__ __forceinline__ void fetch_segment(uint32_t ring, const uint8_t *stream, int seg, int lane) {
    const uint8_t *src = stream + (size_t)seg * kSegBytes + lane * 16;
    const uint32_t dst = ring + (seg & (kRingSegs - 1)) * kSegBytes + lane * 16;
    cp_async16(dst, src);
    if (lane < (kSegBytes - 512) / 16) cp_async16(dst + 512, src + 512);
    asm volatile("cp.async.commit_group;" ::: "memory");
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
#ifndef EXP_NOLDS
            "@p ld.shared.v4.f32 {%4, %5, %6, %7}, [a];\n"
#endif
            "@p fma.rn.f32 %0, %11, %4, %0;\n"
#ifndef EXP_NOFMA
            "@p fma.rn.f32 %1, %11, %5, %1;\n"
            "@p fma.rn.f32 %2, %11, %6, %2;\n"
            "@p fma.rn.f32 %3, %11, %7, %3;\n"
#endif
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
                          int MT, int ntiles) {
    __shared__ float tile[32][33];
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the SpMM kernel may start its prologue
    const int r0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int ii = ty; ii < 32; ii += 8) {
        const int i = i0 + ii, r = r0 + tx;
        tile[ii][tx] = (i < m && r < k) ? a[(size_t)i * lda + r] : 0.0f;
    }
    __syncthreads();
    const int i = i0 + tx;                              // dense column; MT is 16 or a multiple of 32
    const int t = i / MT, il = i - t * MT;
    if (t >= ntiles) return;
    for (int rr = ty; rr < 32; rr += 8) {
        const int r = r0 + rr;
        if (r < k) xt[((size_t)t * k_alloc + r) * MT + il] = tile[tx][rr];
    }
}


// blockDim.x = (nwarps + 1) * 32: nwarps consumer warps and one producer warp (the last).
// RS = W rows per slot (a warp owns 4*RS rows).  SPLIT = 2 | 4 (with NF4 = 1, for m <= 16 | 8): the m-tile is
// 16 | 8 columns wide; the 8 lanes of a slot form SPLIT parts that take SPLIT consecutive entries
// of the slot's row at the same time, and their sums are added at the end -- 1/SPLIT of the
// shared-memory wavefronts per entry, at the price of a summation order that is no longer strictly
// column by column.
template <int RS, int NF4, int STAGES, bool SCALE, int SPLIT>
__global__ void __maxnreg__((RS == 1 && NF4 == 1) ? 64 : TILE_MAXNREG_BIG)
spmm_tile_kernel(const uint8_t *__restrict__ blocks, const int32_t *__restrict__ grp_ptr,
                 const uint32_t *__restrict__ grp_nq, const int32_t *__restrict__ row_ptr, int ngroups, int nchunks,
                 int n, const float *__restrict__ xt, int k_alloc, float *__restrict__ c, int ldc, int m, float alpha,
                 float beta, int nwarps, int epi) {
    constexpr int MT = 32 * NF4 / SPLIT, G = 4 * RS;
    static_assert(SPLIT == 1 || ((SPLIT == 2 || SPLIT == 4) && NF4 == 1), "narrow tiles have one float4 per lane");
    constexpr int STAGE_BYTES = 256 * MT * 4;                    // 255 data rows, padded to a round size
    constexpr uint32_t CHUNK_BYTES = kTileK * MT * 4;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint8_t *ring_base = smem_raw + (size_t)STAGES * STAGE_BYTES;                           // [nwarps][kRingBytes]
    uint64_t *full = reinterpret_cast<uint64_t *>(ring_base + (size_t)nwarps * kRingBytes);  // [STAGES]
    uint64_t *empty = full + STAGES;                                                         // [STAGES]
    uint32_t *s_nq = reinterpret_cast<uint32_t *>(empty + STAGES);                           // [nwarps][nchunks]
    __shared__ int s_len[kMaxRows];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = lane >> 3, c8 = lane & (8 / SPLIT - 1), part = (lane & 7) / (8 / SPLIT);   // column lane, part of the slot
    const int RB = nwarps * G;
    // 1-D grid, m-tile fastest: the CTAs that share a block of W rows are neighbours in launch order,
    // so the quad rows are read from HBM once and served from L2 to the other m-tiles
    const int mtiles = (m + MT - 1) / MT;
    const int row_block = blockIdx.x / mtiles, m_tile = blockIdx.x - row_block * mtiles;
    const int j0 = row_block * RB;
    const int m0 = m_tile * MT;
    const float *xt_tile = xt + (size_t)m_tile * k_alloc * MT;

    TL(0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], nwarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Cold-start loads that nothing depends on yet: the warp's stream position, the quad-row counts
    // of every (warp, chunk) of this CTA (one contiguous piece of grp_nq), the row lengths of the
    // epilogue (kept in a register until then) and L2 prefetches of the C tile updated at the end.
    const int g = row_block * nwarps + warp;
    const bool live = warp < nwarps && g < ngroups;
    const int64_t s0 = live ? grp_ptr[(int64_t)g * nchunks] : 0;
    int my_len = 0;
    if (threadIdx.x < RB && j0 + threadIdx.x < n) my_len = row_ptr[j0 + threadIdx.x + 1] - row_ptr[j0 + threadIdx.x];
    {
        const int64_t first = (int64_t)row_block * nwarps * nchunks, limit = (int64_t)ngroups * nchunks;
        for (int idx = threadIdx.x; idx < nwarps * nchunks; idx += blockDim.x)
            s_nq[idx] = first + idx < limit ? grp_nq[first + idx] : 0u;
        const int lines = (RB + 31) / 32 + 1;                    // 128-byte lines of one C row segment
        for (int idx = threadIdx.x; idx < MT * lines && !(epi & kEpiNoPrefetch); idx += blockDim.x) {
            const int i = idx / lines, l = idx - i * lines;
            if (m0 + i < m && j0 + l * 32 < n)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(c + (size_t)(m0 + i) * ldc + j0 + l * 32));
        }
    }
    __syncthreads();
    TL(1);

    float4 acc[RS][NF4];
#pragma unroll
    for (int rr = 0; rr < RS; ++rr)
#pragma unroll
        for (int f = 0; f < NF4; ++f) acc[rr][f] = make_float4(0.f, 0.f, 0.f, 0.f);

    if (warp == nwarps) {
        // ---- producer: keep the X ring full ----
        if (lane == 0) {
            asm volatile("griddepcontrol.wait;" ::: "memory");   // X^T is written by the kernel before this one
            for (int ch = 0; ch < nchunks; ++ch) {
                const int stage = ch % STAGES, round = ch / STAGES;
                if (round > 0) mbar_wait(&empty[stage], (uint32_t)((round - 1) & 1));
#ifdef EXP_NOTMA
                mbar_arrive(&full[stage]);
#else
                mbar_expect_tx(&full[stage], CHUNK_BYTES);
                bulk_g2s(smem_raw + (size_t)stage * STAGE_BYTES, xt_tile + (size_t)ch * kTileK * MT, CHUNK_BYTES,
                         &full[stage]);
#endif
            }
        }
    } else {
        // ---- consumers ----
        const uint8_t *ring_ptr = ring_base + (size_t)warp * kRingBytes;
        const uint32_t ring = smem_u32(ring_ptr);
        const uint8_t *stream = blocks + s0 * kBlockBytes;        // dead warps read (and ignore) the head of the array
        fetch_segment(ring, stream, 0, lane);
        fetch_segment(ring, stream, 1, lane);
        fetch_segment(ring, stream, 2, lane);
        fetch_segment(ring, stream, 3, lane);
        // One quad row is always held in registers ahead of the one being used.  `seg` is the segment
        // the look-ahead read is in, `left` the quad rows not yet read from it.  Reading past the end
        // of the warp's stream is harmless: the array is padded and the rows are never used.
        const uint8_t *rp = ring_ptr + slot * 16;                // this slot's read pointer into the ring
        const int roff = 64 - slot * 12;                         // from the slot's weights to its offsets
        asm volatile("cp.async.wait_group 3;" ::: "memory");
        __syncwarp();
        TL(2);
        float4 cur_w = *reinterpret_cast<const float4 *>(rp);
        uint32_t cur_r = *reinterpret_cast<const uint32_t *>(rp + roff);
        rp += kBlockBytes;
        int seg = 0, left = kSegRows - 1;
        const uint32_t xlane = smem_u32(smem_raw) + c8 * 16;
        const uint32_t *my_nq = s_nq + warp * nchunks;

        float4 xr[4][NF4];                                       // X staging registers, one set per step
#pragma unroll
        for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int f = 0; f < NF4; ++f) xr[t][f] = make_float4(0.f, 0.f, 0.f, 0.f);
        // the quad rows of one slot-row set in the current chunk
        auto walk = [&](float4 (&a)[NF4], int nq, uint32_t xbase) {
#pragma unroll 2
            for (int q = 0; q < nq; ++q) {
                float4 w = cur_w;
                const uint32_t r4 = cur_r;
                if (left == 0) {                                 // the look-ahead moves into the next segment
                    asm volatile("cp.async.wait_group 2;" ::: "memory");
                    __syncwarp();                                // it has landed for every lane; the one left behind is free
                    ++seg;
                    fetch_segment(ring, stream, seg + kRingSegs - 1, lane);
                    if ((seg & (kRingSegs - 1)) == 0) rp = ring_ptr + slot * 16;
                    left = kSegRows;
                }
                --left;
                cur_w = *reinterpret_cast<const float4 *>(rp);   // next quad row, ahead of use
                cur_r = *reinterpret_cast<const uint32_t *>(rp + roff);
                rp += kBlockBytes;
                if (SCALE) { w.x *= alpha; w.y *= alpha; w.z *= alpha; w.w *= alpha; }
                if (SPLIT == 4) {                                // entries 0..3, one per quarter of the slot
                    const float wp = part < 2 ? (part ? w.y : w.x) : (part == 2 ? w.z : w.w);
                    slot_step<NF4, MT * 4>(a, xr[0], wp, __byte_perm(r4, 0u, 0x4440u + part), xbase);
                } else if (SPLIT == 2) {                         // entries 0|1, then 2|3, one per half of the slot
                    slot_step<NF4, MT * 4>(a, xr[0], part ? w.y : w.x, __byte_perm(r4, 0u, 0x4440u + part), xbase);
                    slot_step<NF4, MT * 4>(a, xr[1], part ? w.w : w.z, __byte_perm(r4, 0u, 0x4442u + part), xbase);
                } else {
                    slot_step<NF4, MT * 4>(a, xr[0], w.x, __byte_perm(r4, 0u, 0x4440u), xbase);
                    slot_step<NF4, MT * 4>(a, xr[1], w.y, __byte_perm(r4, 0u, 0x4441u), xbase);
                    slot_step<NF4, MT * 4>(a, xr[2], w.z, __byte_perm(r4, 0u, 0x4442u), xbase);
                    slot_step<NF4, MT * 4>(a, xr[3], w.w, __byte_perm(r4, 0u, 0x4443u), xbase);
                }
            }
        };


        for (int ch = 0; ch < nchunks; ++ch) {
            const int stage = ch % STAGES;
#ifdef EXP_NOCOMPUTE
            const uint32_t nq2 = 0u;
#else
            const uint32_t nq2 = my_nq[ch];
#endif
            mbar_wait(&full[stage], (uint32_t)((ch / STAGES) & 1));
            if (ch == 0) TL(3);
            const uint32_t xbase = xlane + stage * STAGE_BYTES;
            walk(acc[0], (int)(nq2 & 0xffffu), xbase);
            if (RS == 2) walk(acc[RS - 1], (int)(nq2 >> 16), xbase);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);           // this warp is done with the stage
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");     // nothing of the ring is in flight any more
        TL(4);
    }

    // every warp is past the k loop: the X ring is free, stage the RB x MT result tile in it
    if (threadIdx.x < RB) s_len[threadIdx.x] = my_len;
    // C[m0+i][j0 .. j0+RB) += tile column i: 128-byte rows along j.  A warp takes kEpi rows of C at a
    // time and issues all their loads before the first use; the first (normally the only) batch is
    // issued here, ahead of the two barriers and the staging of the tile, which hide its latency.
    constexpr int kEpi = NF4 == 2 ? 5 : 4, kJ = kMaxRows / 32;
    const int nw_all = blockDim.x >> 5;
    const bool red_add = (epi & kEpiRedAdd) && beta == 1.0f;
    float c0[kEpi][kJ];
    auto load_c = [&](int ib) {
#pragma unroll
        for (int e = 0; e < kEpi; ++e)
#pragma unroll
            for (int u = 0; u < kJ; ++u) {
                const int jl = lane + 32 * u, i = ib + e;
                c0[e][u] = (i < MT && m0 + i < m && jl < RB && j0 + jl < n) ? c[(size_t)(m0 + i) * ldc + j0 + jl] : 0.0f;
            }
    };
    if (!red_add) load_c(warp * kEpi);
    __syncthreads();
    TL(5);
    float *tile = reinterpret_cast<float *>(smem_raw);           // [RB][MT + 1]
    if (warp < nwarps) {
#pragma unroll
        for (int rr = 0; rr < RS; ++rr)
#pragma unroll
            for (int f = 0; f < NF4; ++f) {
                if (SPLIT > 1) {                                 // add the parts' sums (fixed order)
#pragma unroll
                    for (int o = 8 / SPLIT; o < 8; o <<= 1) {
                        acc[rr][f].x += __shfl_xor_sync(0xffffffffu, acc[rr][f].x, o);
                        acc[rr][f].y += __shfl_xor_sync(0xffffffffu, acc[rr][f].y, o);
                        acc[rr][f].z += __shfl_xor_sync(0xffffffffu, acc[rr][f].z, o);
                        acc[rr][f].w += __shfl_xor_sync(0xffffffffu, acc[rr][f].w, o);
                    }
                    if (part) continue;
                }
                float *p = tile + (size_t)(warp * G + rr * 4 + slot) * (MT + 1) + 32 * f + c8 * 4;
                p[0] = acc[rr][f].x; p[1] = acc[rr][f].y; p[2] = acc[rr][f].z; p[3] = acc[rr][f].w;
            }
    }
    __syncthreads();
    if (red_add) {
        // C in a peer GPU's memory: C += sum as a reduction at the memory (red.global.add.f32, one
        // rounding like the add it replaces), so no read of C crosses the link and nothing waits for one
        for (int i = warp; i < MT && m0 + i < m; i += nw_all)
#pragma unroll
            for (int u = 0; u < kJ; ++u) {
                const int jl = lane + 32 * u;
                if (jl < RB && j0 + jl < n && s_len[jl])
                    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(c + (size_t)(m0 + i) * ldc + j0 + jl),
                                 "f"(tile[(size_t)jl * (MT + 1) + i]) : "memory");
            }
        return;
    }
    for (int ib = warp * kEpi; ib < MT && m0 + ib < m; ib += nw_all * kEpi) {
        if (ib != warp * kEpi) load_c(ib);
#pragma unroll
        for (int e = 0; e < kEpi; ++e)
#pragma unroll
            for (int u = 0; u < kJ; ++u) {
                const int jl = lane + 32 * u, i = ib + e;
                if (i >= MT || m0 + i >= m || jl >= RB || j0 + jl >= n) continue;
                const int len = s_len[jl];
                if (len == 0 && beta == 1.0f) continue;          // S column without entries, no scale: untouched
                float v = c0[e][u];
                if (beta != 1.0f) v *= beta;
                c[(size_t)(m0 + i) * ldc + j0 + jl] = len ? v + tile[(size_t)jl * (MT + 1) + i] : v;
            }
    }
    TL(6);
}

#ifdef EXP_TIMELINE
static int g_tl_ctas = 0;
static void dump_timeline() {
    static unsigned long long h[2048 * 8];
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(h, g_timeline, sizeof(h));
    const int n = std::min(g_tl_ctas, 2048);
    unsigned long long t0 = ~0ull;
    for (int i = 0; i < n; ++i) t0 = std::min(t0, h[i * 8]);
    double avg[7] = {0}, mx[7] = {0};
    for (int i = 0; i < n; ++i)
        for (int p = 0; p < 7; ++p) { const double d = (double)(h[i * 8 + p] - t0) * 1e-3; avg[p] += d / n; mx[p] = std::max(mx[p], d); }
    fprintf(stderr, "timeline (us since first CTA start; avg / max over %d CTAs of the last launch):\n", n);
    const char *names[7] = {"start", "prologue done", "stream ring ready", "first X chunk landed", "k loop done (warp 0)", "all warps done", "end"};
    for (int p = 0; p < 7; ++p) fprintf(stderr, "  %-22s %8.2f %8.2f\n", names[p], avg[p], mx[p]);
}
#endif

template <int RS, int NF4, int STAGES, int SPLIT = 1>
static int launch_tile(const sblas_b200_matrix *mat, const TileFormat &tf, const float *xt, int k_alloc, float *d_c,
                       int ldc, int m, float alpha, float beta, int nwarps, int epi, cudaStream_t stream) {
    constexpr int MT = 32 * NF4 / SPLIT, MAXW = kMaxRows / (4 * RS);
    const int nchunks = tf.nchunks;
    auto smem_for = [nchunks](int nw) { return (size_t)STAGES * 256 * MT * 4 + (size_t)nw * kRingBytes + 2 * STAGES * 8 + (size_t)nw * nchunks * 4; };
    if (smem_for(nwarps) > 226 * 1024) return 1;                  // k too long for the per-chunk counts: gather kernel
    static bool configured_on[64] = {false};                     // per template instance and device (the attribute is per context)
    bool &configured = configured_on[mat->device & 63];
    if (!configured) {
        (void)MAXW;
        SB_CUDA(cudaFuncSetAttribute(spmm_tile_kernel<RS, NF4, STAGES, false, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        SB_CUDA(cudaFuncSetAttribute(spmm_tile_kernel<RS, NF4, STAGES, true, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        configured = true;
    }
    const dim3 grid(((tf.ngroups + nwarps - 1) / nwarps) * ((m + MT - 1) / MT));
    {
        char name[64];
        snprintf(name, sizeof(name), "spmm_tile<%d,%d,%d,%d>x%d%s", RS, NF4, STAGES, SPLIT, nwarps, (epi & kEpiRedAdd) ? " red" : "");
        note_product(const_cast<sblas_b200_matrix *>(mat), name, m, 5.0, (int64_t)4 * (mat->n + 1) + (int64_t)8 * tf.ngroups * tf.nchunks);
    }
#ifdef EXP_TIMELINE
    { static bool reg = false; if (!reg) { atexit(dump_timeline); reg = true; } g_tl_ctas = grid.x; }
#endif
    // Programmatic dependent launch: this kernel may start while the transpose before it in the stream
    // is still draining; only the producer warp's X copies wait for it (griddepcontrol.wait).
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3((nwarps + 1) * 32);
    cfg.dynamicSmemBytes = smem_for(nwarps);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const int ngroups = tf.ngroups, n = mat->n;
    auto kernel = alpha == 1.0f ? spmm_tile_kernel<RS, NF4, STAGES, false, SPLIT> : spmm_tile_kernel<RS, NF4, STAGES, true, SPLIT>;
    SB_CUDA(cudaLaunchKernelEx(&cfg, kernel, (const uint8_t *)tf.d_blocks, (const int32_t *)tf.d_grp_ptr,
                               (const uint32_t *)tf.d_grp_nq, (const int32_t *)mat->d_row_ptr, ngroups, nchunks, n, xt,
                               k_alloc, d_c, ldc, m, alpha, beta, nwarps, epi));
    g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

namespace {
struct TilePlan { int rs, nf4, nwarps, split; };

// Pick rows per slot, float4s per lane and consumer warps per CTA.  One CTA per SM at a time, so a
// launch costs (waves of CTAs) x (time of one CTA); the time of one CTA is a fixed part (cold start,
// epilogue) plus quad rows per warp x the time of one quad row, which is bound either by the issue
// slots its warps share (4 schedulers per SM) or by the dependent latency of one iteration.
// Constants fitted to B200 measurements (profiles/): cycles.
TilePlan plan_tile(const sblas_b200_matrix *mat, int m, int sms) {
    const double fixed = 8000.0, eff = 0.75;
    const int nchunks = (mat->k + kTileK - 1) / kTileK;
    const double quads = 1.25 * (double)mat->nnz / std::max(1, mat->n) / 4.0 + 0.5 * nchunks;   // per 4-row set
    // 8- / 16-column tile, 4 / 2 entries of a slot per step; the 8-column tile pays only on large matrices
    // (16384^2 at 90 %: 120 -> 95 us; 4096^2 at 90 %: 26 -> 32 us, profiles/r01c_small_m.md)
    const int split = (m <= 8 && mat->nnz >= (6 << 20)) ? 4 : (m <= 16 ? 2 : 1);
    TilePlan best = {2, 1, 16, split};
    double best_cost = 1e300;
    for (int rs = 1; rs <= 2; ++rs)
        for (int nf4 = 1; nf4 <= 2; ++nf4) {
            if (nf4 == 2 && m <= 32) continue;
            const int max_w = (rs == 1 && nf4 == 1) ? 28 : TILE_MAXW_BIG;
            const double instr = split == 4 ? 16.0 : (split == 2 ? 24.0 : (nf4 == 1 ? 38.0 : 55.0));
            const double lat = split > 1 ? 60.0 : (nf4 == 1 ? 90.0 : 110.0);
            const int mtiles = split > 1 ? 1 : (m + 32 * nf4 - 1) / (32 * nf4);
            const int ngroups = (mat->n + 4 * rs - 1) / (4 * rs);
            for (int nw = 4; nw <= max_w; ++nw) {
                const int64_t ctas = (int64_t)((ngroups + nw - 1) / nw) * mtiles;
                const int64_t waves = (ctas + sms - 1) / sms;
                const double per_quad = std::max(instr * ((nw + 3) / 4) / eff, lat);
                const double cost = (double)waves * (fixed + rs * quads * per_quad);
                if (cost < best_cost * (1.0 - 1e-9)) { best_cost = cost; best = {rs, nf4, nw, split}; }
            }
        }
    return best;
}
}  // namespace

// Variant override for experiments: SBLAS_B200_TILE = "<rows per slot 1|2><float4s per lane 1|2><warps, 2 digits>".
int run_spmm_tile(sblas_b200_matrix *mat, const float *d_a, int m, int lda, float *d_c, int ldc, float alpha,
                  float beta, cudaStream_t stream) {
    static int sms = 0;
    if (!sms) {
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mat->device) != cudaSuccess || sms <= 0) sms = 148;
    }
    TilePlan plan = plan_tile(mat, m, sms);
    if (const char *env = getenv("SBLAS_B200_TILE")) {
        if ((env[0] == '1' || env[0] == '2') && (env[1] == '1' || env[1] == '2') && env[2] && env[3]) {
            plan.rs = env[0] - '0';
            plan.nf4 = plan.split > 1 ? 1 : env[1] - '0';
            plan.nwarps = std::min((plan.rs == 1 && plan.nf4 == 1) ? 28 : TILE_MAXW_BIG, std::max(1, atoi(env + 2)));
        }
    }
    TileFormat &tf = mat->tile[plan.rs == 2];
    if (tf.state == 0) {
        const int brc = build_tile_format(mat, 4 * plan.rs, stream);
        if (brc != SBLAS_B200_OK) return brc;
    }
    if (tf.state != 1) return 1;                                  // no tile layout for this matrix: caller falls back to the gather kernel
    const int MT = 32 * plan.nf4 / plan.split;
    const int ntiles = (m + MT - 1) / MT;
    const int k_alloc = tf.nchunks * kTileK;
    int rc = ensure_buffer(&mat->d_ws, &mat->ws_bytes, (size_t)ntiles * k_alloc * MT * 4);
    if (rc != SBLAS_B200_OK) return rc;
    const dim3 tgrid((mat->k + 31) / 32, (ntiles * MT + 31) / 32);
    SB_LAUNCH(transpose_in_tiled_kernel, tgrid, 256, 0, stream, d_a, m, mat->k, lda, mat->d_ws, k_alloc, MT, ntiles);
    SB_CHECK_LAUNCH();
    struct Timer {                                                // event pair around the dominant kernel when profiling
        sblas_b200_matrix *mat; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
        Timer(sblas_b200_matrix *m_, cudaStream_t s_) : mat(m_), st(s_) {
            if (mat->profiling) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, st); }
        }
        ~Timer() { if (a) { cudaEventRecord(b, st); mat->prof_events.push_back(std::make_pair(a, b)); } }
    } timer(mat, stream);
    const float *xt = mat->d_ws;
    // where C lives decides the epilogue: a pointer into another GPU's memory (sblas_b200_peer_open) is
    // neither prefetched nor read when beta == 1.  SBLAS_B200_EPI overrides (experiments).
    static const int env_epi = [] { const char *e = getenv("SBLAS_B200_EPI"); return (e && *e) ? atoi(e) : -1; }();
    int epi = env_epi;
    if (epi < 0) epi = is_peer_pointer(d_c) ? (kEpiNoPrefetch | kEpiRedAdd) : 0;
    if (plan.split == 4 && plan.rs == 1) return launch_tile<1, 1, 4, 4>(mat, tf, xt, k_alloc, d_c, ldc, m, alpha, beta, plan.nwarps, epi, stream);
    if (plan.split == 4) return launch_tile<2, 1, 4, 4>(mat, tf, xt, k_alloc, d_c, ldc, m, alpha, beta, plan.nwarps, epi, stream);
    if (plan.split == 2 && plan.rs == 1) return launch_tile<1, 1, 4, 2>(mat, tf, xt, k_alloc, d_c, ldc, m, alpha, beta, plan.nwarps, epi, stream);
    if (plan.split == 2) return launch_tile<2, 1, 4, 2>(mat, tf, xt, k_alloc, d_c, ldc, m, alpha, beta, plan.nwarps, epi, stream);
    if (plan.rs == 1 && plan.nf4 == 1) return launch_tile<1, 1, 3>(mat, tf, xt, k_alloc, d_c, ldc, m, alpha, beta, plan.nwarps, epi, stream);
    if (plan.rs == 1) return launch_tile<1, 2, 2>(mat, tf, xt, k_alloc, d_c, ldc, m, alpha, beta, plan.nwarps, epi, stream);
    if (plan.nf4 == 1) return launch_tile<2, 1, 3>(mat, tf, xt, k_alloc, d_c, ldc, m, alpha, beta, plan.nwarps, epi, stream);
    return launch_tile<2, 2, 2>(mat, tf, xt, k_alloc, d_c, ldc, m, alpha, beta, plan.nwarps, epi, stream);
}

}  // namespace sblas_b200
