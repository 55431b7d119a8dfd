// spmm_tc.cu -- K-spmm-tc: the product for m >= 32 on matrices dense enough that multiplying the zeros on the
// tensor cores beats gathering the non-zeros on the CUDA cores (profiles/r02_spmm_tc.md: from ~8 % of the elements
// stored; the gather kernels are bound by one 128-byte shared-memory wavefront per non-zero and 32 dense columns,
// i.e. 25 % of the fp32 peak).
//
// Replaces the reference's AXPY micro-kernel (kernel.cc:371-584 driven by :771-800) like spmm_tile.cu does.
// Y[n x m] = W X as a dense contraction: D[128 W rows x N dense columns] += W_tile[128 x 32] * X_tile[N x 32]^T per
// 32-column step of k, on tcgen05 (UTCMMA) with fp32 accumulators in tensor memory.  fp32 accuracy from TF32
// products by splitting both operands: w = w_hi + w_lo, x = x_hi + x_lo (each with 13 low mantissa bits clear, so
// nothing is rounded on the way into the MMA), D += w_hi x_hi + w_hi x_lo + w_lo x_hi; the dropped w_lo x_lo term is
// 2^-22 of the product.  Measured backward error against fp64: ~2e-7, the level of the reference's own arithmetic.
//
//   W tile   never stored as fp32: 4 warps expand it per step from the one-byte dense code matrix of W
//            (spmv_dense.cu builds it: n x k_pad codes, >= T where there is no entry) through the codebook in shared
//            memory (16 copies, split into hi and lo as it goes) into TENSOR MEMORY with tcgen05.st -- the MMA takes
//            its A operand from there (lane = W row, one column per k), so W never crosses shared memory: with both
//            operands in shared memory the TF32 MMAs' own operand reads (12 KB per 136 clocks) plus the fills ran the
//            shared memory out of bandwidth (profiles/r02_spmm_tc.md)
//   X tiles  split and laid out as operand images by tc_split_x_kernel (one pass over A), then brought in by ONE
//            bulk copy per tile and stage (cp.async.bulk -> mbarrier complete_tx, UBLKCP)
//   MMA      one elected thread: per stage 4 k-substeps x 3 products, tcgen05.commit hands the stage back
//   epilogue tcgen05.ld 32 lanes x 32 columns per warp, C = beta*C + D in 128-byte rows along j; rows of W
//            without entries only get the scale (sparse-matrix.cc:57,164), like every other kernel here
//   split k  when the tiles do not fill the machine: partial tiles to a workspace, the last CTA of a tile adds
//            them in split order (fixed order: deterministic)
//
// Semantics kept: absent entries never touch x in the reference, so 0 * Inf must not appear: tc_split_x_kernel
// flags non-finite x, the MMA kernel then leaves C alone and tc_fallback_kernel (normally an empty launch)
// computes the product from the CSR.  Non-finite codebook entries or alpha keep the matrix off this path.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>

#include "async_copy.cuh"
#include "device_matrix.h"

namespace sblas_b200 {

namespace {

constexpr int kTcM = 128;               // W rows per CTA = MMA M = tensor-memory lanes
constexpr int kTcBK = 32;               // k columns per stage: one 128-byte swizzle row of fp32
constexpr int kTcTabCopies = 16;         // codebook copies in shared memory: lanes l and l + 16 share one (two-way conflicts at most)
constexpr int kTcWStages = 4;           // A operands (W_hi | W_lo, 64 columns) kept in tensor memory
constexpr int kTcEWarps = 8;            // warps 0-3 expand the even steps' W operands, warps 4-7 the odd ones; all eight run the epilogue
constexpr int kTcThreads = (kTcEWarps + 2) * 32;   // warp 8 feeds X, warp 9 issues the MMAs

__device__ __forceinline__ float tf32_floor(float v) { return __uint_as_float(__float_as_uint(v) & 0xffffe000u); }

// one lane of a converged warp (elect.sync): the form the compiler turns into straight uniform-datapath instructions --
// behind a plain `lane == 0` test every tcgen05.mma was wrapped in its own elect-and-branch loop
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .b32 rx;\n"
        ".reg .pred px;\n"
        "elect.sync rx|px, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, px;\n"
        "}\n" : "=r"(pred));
    return pred != 0;
}

// bounded wait: a protocol error must end in a trap, not in a hung GPU
__device__ __forceinline__ void mbar_wait_or_trap(uint64_t *bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    for (uint32_t spin = 0;; ++spin) {
        uint32_t ok;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
        if (ok) return;
        if (spin > (1u << 28)) __trap();
    }
}

// shared-memory matrix descriptor: K-major, SWIZZLE_128B, 8-row atoms 1024 bytes apart (cute::UMMA::SmemDescriptor:
// start address >> 4 in bits 0-13, leading byte offset (unused for swizzled K-major, 1) in 16-29, stride byte offset
// >> 4 in 32-45, version 1 in 46-47, layout type 2 = SWIZZLE_128B in 61-63)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}

// A from tensor memory (lane = row, one 32-bit column per k), B from shared memory
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}

__device__ __forceinline__ void umma_tf32_ts_acc(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.eq.b32 p, 0, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc) : "memory");
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, "
        "%19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]),
        "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]),
        "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
}

__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, "
        "%19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// the same 32 columns of every accumulator, added in accumulator order
template <int NACC, int N_TILE>
__device__ __forceinline__ void tmem_ld_sum(uint32_t taddr, float (&v)[32]) {
    tmem_ld32(taddr, v);
#pragma unroll
    for (int a = 1; a < NACC; ++a) {
        float w[32];
        tmem_ld32(taddr + a * N_TILE, w);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += w[i];
    }
}

#ifdef EXP_TC_TIMELINE
// experiment build only (tools/build_variant.sh): phase stamps of every CTA of the last contraction, dumped at exit
__device__ unsigned long long g_tc_timeline[4096 * 16];
__device__ __forceinline__ unsigned long long tc_gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define TCL(p, tid) do { if (threadIdx.x == (tid)) { const unsigned b_ = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x; \
                         if (b_ < 4096) g_tc_timeline[b_ * 16 + (p)] = tc_gtime(); } } while (0)
#else
#define TCL(p, tid) do { } while (0)
#endif

// ---- X -> operand images ----------------------------------------------------------------------------------------
// xh / xl [m tile][k block][N rows x 128 bytes]: row i of a tile = dense row m0 + i (a row of A: contiguous in k),
// its eight 16-byte chunks stored at chunk ^ (i & 7) -- the SWIZZLE_128B image the MMA reads, so the main kernel moves
// a tile with one plain bulk copy.  Rows beyond m and columns beyond k are zeros.  Non-finite values raise the flag.
__global__ void __launch_bounds__(256)
tc_split_x_kernel(const float *__restrict__ a, int m, int k, int lda, int n_tile, int nkb, float *__restrict__ xh,
                  float *__restrict__ xl, uint32_t *__restrict__ flag, uint32_t call_id, int vec) {
    const int kb = blockIdx.x, mt = blockIdx.y;
    const size_t tile = ((size_t)mt * nkb + kb) * n_tile * kTcBK;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the contraction may set up its CTAs meanwhile
    // one (row, 16-byte chunk) per thread: grid.z x 256 threads cover the n_tile x 8 chunks of the tile
    const int t = blockIdx.z * 256 + threadIdx.x;
    if (t >= n_tile * 8) return;
    const int i = t >> 3, ch = t & 7;
    const int gi = mt * n_tile + i, r0 = kb * kTcBK + ch * 4;
    float v[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (gi < m) {
        const float *src = a + (size_t)gi * lda + r0;
        if (vec && r0 + 3 < k) {
            const float4 q = __ldg(reinterpret_cast<const float4 *>(src));
            v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        } else {
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (r0 + u < k) v[u] = src[u];
        }
    }
    float4 hi, lo;
    hi.x = tf32_floor(v[0]); hi.y = tf32_floor(v[1]); hi.z = tf32_floor(v[2]); hi.w = tf32_floor(v[3]);
    lo.x = tf32_floor(v[0] - hi.x); lo.y = tf32_floor(v[1] - hi.y); lo.z = tf32_floor(v[2] - hi.z); lo.w = tf32_floor(v[3] - hi.w);
    const bool bad = !(isfinite(v[0]) && isfinite(v[1]) && isfinite(v[2]) && isfinite(v[3]));
    const size_t off = tile + (size_t)i * kTcBK + ((ch ^ (i & 7)) << 2);
    *reinterpret_cast<float4 *>(xh + off) = hi;
    *reinterpret_cast<float4 *>(xl + off) = lo;
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicExch(flag, call_id);
}

// ---- the contraction ----------------------------------------------------------------------------------------------
// grid (row blocks of 128 W rows, m tiles, k splits).
// Tensor memory (all 512 columns): [0, N_TILE) the fp32 accumulators; from column 256 a ring of kTcWStages A operands,
// 64 columns each: W_hi (32 k columns of the step, one 32-bit column per k) and W_lo.  The expanders write them with
// tcgen05.st (lane = W row), the MMA reads A from tensor memory and B (the X images) from shared memory.
// Dynamic shared memory, 1024-byte aligned by hand: XS x [X_hi N*128 | X_lo N*128], codebook copies, mbarriers.
template <int N_TILE, int XS>
__global__ void __launch_bounds__(kTcThreads, 1)
spmm_tc_kernel(const uint8_t *__restrict__ dense, int64_t kpad, const int32_t *__restrict__ row_ptr,
               const float *__restrict__ table, int T, float alpha, const float *__restrict__ xh,
               const float *__restrict__ xl, int nkb, int n, int m, float *__restrict__ c, int ldc, float beta,
               const uint32_t *__restrict__ flag, uint32_t call_id, int ksplit, float *__restrict__ partial,
               uint32_t *__restrict__ counters, int red, int dbg) {
    constexpr int X_BYTES = N_TILE * 128;
    constexpr int STAGE_BYTES = 2 * X_BYTES;
    constexpr int WS = kTcWStages;
    // one accumulator whatever the tile width: an output element is then formed by the same sequence of MMAs for every m
    // (round-robin over 256 / N_TILE accumulators was measured and bought nothing once the MMAs were issued back to back)
    constexpr int NACC = 1;
    extern __shared__ uint8_t smem_raw[];
    TCL(0, 0);
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const uint32_t sm_addr = smem_u32(sm);
    float *tab_s = reinterpret_cast<float *>(sm + XS * STAGE_BYTES);    // [256][kTcTabCopies]: entry c for lane l at c * 16 + (l & 15)
    uint64_t *full_x = reinterpret_cast<uint64_t *>(tab_s + 256 * kTcTabCopies), *empty_x = full_x + XS, *full_w = empty_x + XS,
             *empty_w = full_w + WS, *acc_full = empty_w + WS;
    uint64_t *fix_bar = acc_full + 1;                           // [2]: partial tiles of other shares have arrived (last CTA of a tile)
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(acc_full + 3);
    uint32_t *last_slot = tmem_slot + 1;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int j0 = blockIdx.x * kTcM, mt = blockIdx.y, m0 = mt * N_TILE;
    // this CTA's share of the k blocks, in groups of four (one 128-byte line of codes per W row)
    const int split = blockIdx.z;
    const int kb_lo = 4 * (int)((int64_t)(nkb / 4) * split / ksplit), kb_hi = 4 * (int)((int64_t)(nkb / 4) * (split + 1) / ksplit);
    const int steps = kb_hi - kb_lo;

    // one global load per thread, in flight during the set-up below (a loop over the 4096 copies was 3.6 us of dependent
    // loads at the head of every CTA, profiles/r02_spmm_tc.md)
    const float tab_mine = (threadIdx.x < 256 && (int)threadIdx.x < T) ? __ldg(table + threadIdx.x) * alpha : 0.0f;   // val_table[code] * alpha (kernel.cc:791); split when expanded
    // the expanders' first line of codes travels meanwhile as well
    const int e_set = warp >> 2, e_r = (warp & 3) * 32 + lane, e_j = j0 + e_r;
    const bool e_have = warp < kTcEWarps && e_j < n && steps > 0 && !(dbg & 1);
    const uint8_t *e_src = dense + (size_t)min(e_j, n - 1) * kpad + (size_t)kb_lo * kTcBK;
    uint4 nxt[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
        nxt[u] = e_have ? __ldg(reinterpret_cast<const uint4 *>(e_src) + 2 * e_set + (u & 1) + 4 * (u >> 1))
                        : make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
    if (threadIdx.x < 256) {
        const float4 q = make_float4(tab_mine, tab_mine, tab_mine, tab_mine);
#pragma unroll
        for (int cpy = 0; cpy < kTcTabCopies / 4; ++cpy) reinterpret_cast<float4 *>(tab_s + threadIdx.x * kTcTabCopies)[cpy] = q;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < XS; ++s) { mbar_init(&full_x[s], 1); mbar_init(&empty_x[s], 1); }
        for (int s = 0; s < WS; ++s) { mbar_init(&full_w[s], 4); mbar_init(&empty_w[s], 1); }
        mbar_init(acc_full, 1);
        mbar_init(&fix_bar[0], 1);
        mbar_init(&fix_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Launched programmatically dependent on tc_split_x_kernel: the set-up and the expansion of the first W operands run
    // while that kernel drains.  What it wrote is read behind a griddepcontrol.wait: the X images by the producer warp,
    // the flag by the epilogue (non-finite x: nothing is stored, tc_fallback_kernel computes the product).
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the fallback kernel may be set up as well
    TCL(1, 0);
    if (warp == kTcEWarps) {                                    // all of tensor memory: one CTA per SM
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;
    const uint32_t tmem_w = tmem + 256u;                        // the A-operand ring
    TCL(2, 0);

    if (warp == kTcEWarps) {
        // ---- X producer (the whole warp walks the loop, one elected lane issues) ----
        const float *xh_t = xh + ((size_t)mt * nkb + kb_lo) * N_TILE * kTcBK, *xl_t = xl + ((size_t)mt * nkb + kb_lo) * N_TILE * kTcBK;
        asm volatile("griddepcontrol.wait;" ::: "memory");     // the X images are complete
        for (int it = 0; it < steps; ++it) {
            const int s = it % XS;
            if (it >= XS) mbar_wait_or_trap(&empty_x[s], (uint32_t)((it / XS - 1) & 1));
            if (elect_one()) {
                uint8_t *dst = sm + s * STAGE_BYTES;
                mbar_expect_tx(&full_x[s], 2 * X_BYTES);
                bulk_g2s(dst, xh_t + (size_t)it * N_TILE * kTcBK, X_BYTES, &full_x[s]);
                bulk_g2s(dst + X_BYTES, xl_t + (size_t)it * N_TILE * kTcBK, X_BYTES, &full_x[s]);
            }
            __syncwarp();
        }
    } else if (warp == kTcEWarps + 1) {
        // ---- MMA issuer (the whole warp walks the loop, one elected lane issues) ----
        // instruction descriptor (cute::UMMA::InstrDescriptor): D fp32 (bits 4-5 = 1), A and B TF32 (7-9, 10-12 = 2),
        // both K-major (15, 16 = 0), N >> 3 in 17-22, M >> 4 in 24-28
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N_TILE >> 3) << 17) | ((uint32_t)(kTcM >> 4) << 24);
        const uint64_t desc0 = umma_desc(sm_addr);              // the descriptor of shared-memory offset 0: offsets are added to its low bits
        int sx = 0, sw = 0;
        uint32_t px = 0, pw = 0;
        for (int it = 0; it < steps; ++it) {
            mbar_wait_or_trap(&full_w[sw], pw);
            mbar_wait_or_trap(&full_x[sx], px);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (it == 0) TCL(3, (kTcEWarps + 1) * 32);
            if (elect_one()) {
                const uint64_t b_hi = desc0 + (uint64_t)((sx * STAGE_BYTES) >> 4), b_lo = b_hi + (X_BYTES >> 4);
                const uint32_t w_hi = tmem_w + (uint32_t)sw * 64u, w_lo = w_hi + 32u;
                // K = 8 per MMA: 8 columns of A, 32 bytes along the swizzled rows of B.  The twelve MMAs of a step go round
                // NACC accumulators (independent chains for narrow tiles); the epilogue adds the accumulators.
#pragma unroll
                for (int ks = 0; ks < kTcBK / 8; ++ks) {
#pragma unroll
                    for (int pr = 0; pr < 3; ++pr) {
                        const int idx = ks * 3 + pr;
                        const uint32_t d = tmem + (uint32_t)((idx % NACC) * N_TILE);
                        const uint32_t a = (pr == 2 ? w_lo : w_hi) + ks * 8;
                        const uint64_t b = (pr == 1 ? b_lo : b_hi) + 2 * ks;
                        if (it == 0 && idx < NACC) umma_tf32_ts(d, a, b, idesc, 0u);
                        else umma_tf32_ts_acc(d, a, b, idesc);
                    }
                }
                umma_commit(&empty_x[sx]);                      // both rings' slots are free once these MMAs have read them
                umma_commit(&empty_w[sw]);
            }
            __syncwarp();
            if (++sx == XS) { sx = 0; px ^= 1u; }
            if (++sw == WS) { sw = 0; pw ^= 1u; }
        }
        if (elect_one()) umma_commit(acc_full);
        __syncwarp();
    } else {
        // ---- W expanders (one thread per W row of the block = its tensor-memory lane; two sets of four warps take
        //      alternate steps), then the epilogue ----
        const int set = e_set, quad = warp & 3;
        const int r = e_r, j = e_j;
        const uint8_t *src = e_src;
        const bool have = j < n;
        const float *tab_l = tab_s + (lane & (kTcTabCopies - 1));
        const uint32_t t_row = tmem_w + ((uint32_t)(quad * 32) << 16);
        // the codes of four steps are one 128-byte line of the row; this set uses steps set and set + 2 of it (32 bytes each).
        // The next line travels while this one is expanded (the first was requested at the head of the kernel).
        for (int it0 = 0; it0 < steps; it0 += 4) {
            uint4 cur[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) cur[u] = nxt[u];
            if (have && it0 + 4 < steps && !(dbg & 1)) {
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    nxt[u] = __ldg(reinterpret_cast<const uint4 *>(src + (size_t)(it0 + 4) * kTcBK) + 2 * set + (u & 1) + 4 * (u >> 1));
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int it = it0 + set + 2 * q, sw = it % WS;
                if (it >= WS) mbar_wait_or_trap(&empty_w[sw], (uint32_t)((it / WS - 1) & 1));
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t words[8] = {cur[2 * q].x, cur[2 * q].y, cur[2 * q].z, cur[2 * q].w,
                                           cur[2 * q + 1].x, cur[2 * q + 1].y, cur[2 * q + 1].z, cur[2 * q + 1].w};
                uint32_t hi[32], lo[32];
#pragma unroll
                for (int ch = 0; ch < 8; ++ch) {                // four codes of the row at a time
                    const uint32_t wd = words[ch];
                    const float v0 = tab_l[(wd << 4) & 0xff0u], v1 = tab_l[(wd >> 4) & 0xff0u], v2 = tab_l[(wd >> 12) & 0xff0u],
                                v3 = tab_l[(wd >> 20) & 0xff0u];
                    const float h0 = tf32_floor(v0), h1 = tf32_floor(v1), h2 = tf32_floor(v2), h3 = tf32_floor(v3);
                    hi[4 * ch] = __float_as_uint(h0); hi[4 * ch + 1] = __float_as_uint(h1);
                    hi[4 * ch + 2] = __float_as_uint(h2); hi[4 * ch + 3] = __float_as_uint(h3);
                    lo[4 * ch] = __float_as_uint(tf32_floor(v0 - h0)); lo[4 * ch + 1] = __float_as_uint(tf32_floor(v1 - h1));
                    lo[4 * ch + 2] = __float_as_uint(tf32_floor(v2 - h2)); lo[4 * ch + 3] = __float_as_uint(tf32_floor(v3 - h3));
                }
                tmem_st32(t_row + (uint32_t)sw * 64u, hi);
                tmem_st32(t_row + (uint32_t)sw * 64u + 32u, lo);
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&full_w[sw]);
            }
        }

        // epilogue: lane = W row j (tensor-memory lane 32 * quad + lane), 32 dense columns per tcgen05.ld; the two warps of a
        // lane quadrant take alternate 32-column chunks
        const int len = have ? row_ptr[j + 1] - row_ptr[j] : 0;
        if (steps > 0) mbar_wait_or_trap(acc_full, 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        TCL(4, 0);
        asm volatile("griddepcontrol.wait;" ::: "memory");     // (long since complete) the flag below is the split kernel's
        const bool flagged = __ldcg(flag) == call_id;
        const uint32_t t_lane = tmem + ((uint32_t)(quad * 32) << 16);
        const bool touch = have && !(len == 0 && beta == 1.0f);        // S column without entries, no scale: untouched
        if (flagged) {
            // non-finite x: the accumulators hold NaNs, nothing leaves this kernel
        } else if (ksplit == 1) {
            for (int col0 = 32 * set; col0 < N_TILE && m0 + col0 < m; col0 += 64) {
                float acc[32], c_old[32];
                tmem_ld_sum<NACC, N_TILE>(t_lane + col0, acc);
                if (red) {                                      // C in a peer GPU's memory, beta == 1: add at the owner, never read over the link
#pragma unroll
                    for (int t = 0; t < 32; ++t)
                        if (have && len && m0 + col0 + t < m)
                            asm volatile("red.global.add.f32 [%0], %1;" ::"l"(c + (size_t)(m0 + col0 + t) * ldc + j), "f"(acc[t]) : "memory");
                    continue;
                }
#pragma unroll
                for (int t = 0; t < 32; ++t) c_old[t] = (touch && m0 + col0 + t < m) ? c[(size_t)(m0 + col0 + t) * ldc + j] : 0.0f;
#pragma unroll
                for (int t = 0; t < 32; ++t) {
                    if (!touch || m0 + col0 + t >= m) continue;
                    float v = c_old[t];
                    if (beta != 1.0f) v *= beta;
                    c[(size_t)(m0 + col0 + t) * ldc + j] = len ? v + acc[t] : v;   // no entry: exactly the scaled value (keeps -0)
                }
            }
        } else {
            // partial tile [N_TILE][128] of this share; the last CTA of a tile to get here adds the shares in share order.
            // counters == nullptr: the whole grid is one wave of CTAs and tc_fixup_kernel adds the shares with every SM (here
            // the last CTA of each tile would do it alone, all of them at the same time at the end of the only wave)
            const size_t tile_id = (size_t)blockIdx.y * gridDim.x + blockIdx.x;
            float *mine = partial + (tile_id * ksplit + split) * (size_t)N_TILE * kTcM;
            // every other share already there (the usual case in grids of several waves: a tile's shares run in different
            // waves): this CTA is the last one and keeps its share in tensor memory instead of writing it out
            // (only the highest share looks: it is launched last, and the look costs an L2 round trip)
            bool early = false;
            if (counters != nullptr && split == ksplit - 1) {
                if (threadIdx.x == 0) {
                    uint32_t seen;
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counters + tile_id) : "memory");
                    *last_slot = seen == (uint32_t)(ksplit - 1) ? 2u : 0u;
                }
                asm volatile("bar.sync 1, 256;" ::: "memory");  // the eight epilogue warps
                early = *last_slot == 2u;
            }
            if (!early) {
                for (int col0 = 32 * set; col0 < N_TILE && m0 + col0 < m; col0 += 64) {
                    float acc[32];
                    if (steps > 0) tmem_ld_sum<NACC, N_TILE>(t_lane + col0, acc);
                    else {
#pragma unroll
                        for (int t = 0; t < 32; ++t) acc[t] = 0.0f;
                    }
#pragma unroll
                    for (int t = 0; t < 32; ++t) mine[(size_t)(col0 + t) * kTcM + r] = acc[t];
                }
                if (counters != nullptr) {
                    __threadfence();
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    if (threadIdx.x == 0) *last_slot = atomicAdd(&counters[tile_id], 1u) == (uint32_t)(ksplit - 1) ? 1u : 0u;
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                }
            }
            TCL(5, 0);
            if (counters != nullptr && *last_slot) {
                // The other shares' partial tiles come in by bulk copy, 64 dense columns (32 KB per share) at a time, into the
                // idle X ring -- every MMA has completed, so has every copy of the X producer -- and are added from shared
                // memory, this CTA's own share from tensor memory at its place in the order (share 0, 1, 2, ...: the same sum
                // whichever CTA comes last).  Rounds of (column group, HS shares) alternate between the two halves of the
                // ring: the copies of round i + 1 travel while round i is added; the C tile is prefetched into L2 up front.
                // (Plain loads, 80 in flight per thread, took 27 us per tile of N = 256 here and the same copies one round at a
                // time 32 us: every step of a round is a latency of its own, profiles/r02_spmm_tc.md.)
                __threadfence();
                TCL(8, 0);
                constexpr int GC = 64, SLOT = GC * kTcM * 4, HS = XS * STAGE_BYTES / SLOT / 2;
                static_assert(HS >= 1, "each half of the X ring holds at least one 32 KB slot");
                const float *all = partial + tile_id * ksplit * (size_t)N_TILE * kTcM;
                if (!red && (ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(c) & 15) == 0) {
                    const int t = threadIdx.x;                  // dense column m0 + t: 512 contiguous bytes of C
                    if (t < N_TILE && m0 + t < m && j0 + kTcM <= n)
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], 512;" ::"l"(c + (size_t)(m0 + t) * ldc + j0) : "memory");
                }
                // a round = one column group x up to HS of the other shares (o = 0 .. ksplit - 2, share o or o + 1 past this
                // CTA's own); the own share is added in the round its place in the order falls into
                const int nothers = ksplit - 1, nbatch = (nothers + HS - 1) / HS;
                const int ngroups = (min(N_TILE, m - m0) + GC - 1) / GC, rounds = ngroups * nbatch;
                auto share_of = [&](int o) { return o < split ? o : o + 1; };
                auto issue = [&](int rd) {                      // thread 0: the copies of round rd into half rd & 1
                    const int g0 = (rd / nbatch) * GC, o_lo = (rd % nbatch) * HS, o_hi = min(o_lo + HS, nothers);
                    uint64_t *bar = &fix_bar[rd & 1];
                    uint8_t *half = sm + (size_t)(rd & 1) * HS * SLOT;
                    mbar_expect_tx(bar, (uint32_t)(o_hi - o_lo) * SLOT);
                    for (int o = o_lo; o < o_hi; ++o)
                        bulk_g2s(half + (o - o_lo) * SLOT, all + ((size_t)share_of(o) * N_TILE + g0) * kTcM, SLOT, bar);
                };
                TCL(13, 0);
                if (threadIdx.x == 0) {
                    asm volatile("fence.proxy.async;" ::: "memory");   // written with ordinary stores, read by the bulk copies
                    TCL(14, 0);
                    issue(0);
                }
                TCL(15, 0);
                uint32_t used0 = 0u, used1 = 0u;                // completed phases of the two barriers
                float sum[32], c_old[32];
                for (int rd = 0; rd < rounds; ++rd) {
                    const int g0 = (rd / nbatch) * GC, b = rd % nbatch, o_lo = b * HS, o_hi = min(o_lo + HS, nothers);
                    const int sb = b == 0 ? 0 : share_of(o_lo), se = b == nbatch - 1 ? ksplit : share_of(o_hi);   // shares sb .. se - 1
                    const int col0 = g0 + 32 * set;              // this thread: row r, columns col0 .. col0 + 31
                    if (threadIdx.x == 0 && rd + 1 < rounds) issue(rd + 1);   // its half was released by the barrier that ended round rd - 1
                    if (b == 0) {
#pragma unroll
                        for (int t = 0; t < 32; ++t)
                            c_old[t] = (touch && !red && m0 + col0 + t < m) ? c[(size_t)(m0 + col0 + t) * ldc + j] : 0.0f;
                        if (rd == 0) TCL(9, 0);
                    }
                    if (rd & 1) { mbar_wait_or_trap(&fix_bar[1], used1 & 1u); ++used1; }
                    else { mbar_wait_or_trap(&fix_bar[0], used0 & 1u); ++used0; }
                    if (rd == 0) TCL(10, 0);
                    const uint8_t *half = sm + (size_t)(rd & 1) * HS * SLOT;
                    for (int sp = sb; sp < se; ++sp) {
                        float v[32];
                        if (sp == split) {
                            if (steps > 0) tmem_ld_sum<NACC, N_TILE>(t_lane + col0, v);
                            else {
#pragma unroll
                                for (int t = 0; t < 32; ++t) v[t] = 0.0f;
                            }
                        } else {
                            const float *sl = reinterpret_cast<const float *>(half + ((sp < split ? sp : sp - 1) - o_lo) * SLOT) + (size_t)(32 * set) * kTcM + r;
#pragma unroll
                            for (int t = 0; t < 32; ++t) v[t] = sl[t * kTcM];
                        }
                        if (sp == 0) {
#pragma unroll
                            for (int t = 0; t < 32; ++t) sum[t] = v[t];
                        } else {
#pragma unroll
                            for (int t = 0; t < 32; ++t) sum[t] += v[t];
                        }
                    }
                    if (rd == 0) TCL(11, 0);
                    asm volatile("bar.sync 1, 256;" ::: "memory");   // this half may be overwritten (by the copies of round rd + 2)
                    if (b == nbatch - 1) {
#pragma unroll
                        for (int t = 0; t < 32; ++t) {
                            if (!touch || m0 + col0 + t >= m) continue;
                            if (red) {
                                if (len) asm volatile("red.global.add.f32 [%0], %1;" ::"l"(c + (size_t)(m0 + col0 + t) * ldc + j), "f"(sum[t]) : "memory");
                                continue;
                            }
                            float cv = c_old[t];
                            if (beta != 1.0f) cv *= beta;
                            c[(size_t)(m0 + col0 + t) * ldc + j] = len ? cv + sum[t] : cv;
                        }
                        if (rd == nbatch - 1) TCL(12, 0);
                    }
                }
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (threadIdx.x == 0) counters[tile_id] = 0u;   // ready for the next product
            }
        }
        TCL(6, 0);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    TCL(7, 0);
    if (warp == kTcEWarps) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
    }
}

// ---- the shares of k added by a kernel of their own (grids of one wave) -------------------------------------------
// grid (row blocks, m tiles, groups of 8 dense columns), 128 threads = the W rows of the block.  Same sums in the same
// order as the last CTA of a tile forms them in spmm_tc_kernel.
__global__ void __launch_bounds__(128)
tc_fixup_kernel(const float *__restrict__ partial, int ksplit, int n_tile, const int32_t *__restrict__ row_ptr, int n, int m,
                float *__restrict__ c, int ldc, float beta, int red, const uint32_t *__restrict__ flag, uint32_t call_id) {
    constexpr int CU = 8, SG = 3;
    asm volatile("griddepcontrol.wait;" ::: "memory");         // launched programmatically dependent on the contraction
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (__ldcg(flag) == call_id) return;
    const int r = threadIdx.x, j = blockIdx.x * kTcM + r, m0 = blockIdx.y * n_tile, col0 = blockIdx.z * CU;
    if (m0 + col0 >= m) return;
    const size_t tile_id = (size_t)blockIdx.y * gridDim.x + blockIdx.x;
    const float *all = partial + tile_id * ksplit * (size_t)n_tile * kTcM;
    const bool have = j < n;
    const int len = have ? row_ptr[j + 1] - row_ptr[j] : 0;
    const bool touch = have && !(len == 0 && beta == 1.0f);
    float sum[CU], v[CU];
#pragma unroll
    for (int u = 0; u < CU; ++u) {
        sum[u] = __ldcg(all + (size_t)(col0 + u) * kTcM + r);
        v[u] = (touch && !red && m0 + col0 + u < m) ? c[(size_t)(m0 + col0 + u) * ldc + j] : 0.0f;
    }
    for (int sp = 1; sp < ksplit; sp += SG) {
        float p[SG][CU];
#pragma unroll
        for (int g = 0; g < SG; ++g) {
            const float *src = all + ((size_t)min(sp + g, ksplit - 1) * n_tile + col0) * kTcM + r;
#pragma unroll
            for (int u = 0; u < CU; ++u) p[g][u] = __ldcg(src + (size_t)u * kTcM);
        }
#pragma unroll
        for (int g = 0; g < SG; ++g) {
            const bool use = sp + g < ksplit;
#pragma unroll
            for (int u = 0; u < CU; ++u) sum[u] = use ? sum[u] + p[g][u] : sum[u];
        }
    }
#pragma unroll
    for (int u = 0; u < CU; ++u) {
        if (!touch || m0 + col0 + u >= m) continue;
        if (red) {
            if (len) asm volatile("red.global.add.f32 [%0], %1;" ::"l"(c + (size_t)(m0 + col0 + u) * ldc + j), "f"(sum[u]) : "memory");
            continue;
        }
        float cv = v[u];
        if (beta != 1.0f) cv *= beta;
        c[(size_t)(m0 + col0 + u) * ldc + j] = len ? cv + sum[u] : cv;
    }
}

// ---- the product when x is not finite: plain CSR walk, any speed --------------------------------------------------
// Normally an empty launch (flag != call_id).  Warp per W row, lanes over 32 dense rows at a time; weights and order
// as in the gather kernels (alpha folded into the weight, terms added in column order).
__global__ void __launch_bounds__(256)
tc_fallback_kernel(const uint32_t *__restrict__ flag, uint32_t call_id, const int32_t *__restrict__ row_ptr,
                   const int32_t *__restrict__ col, const uint8_t *__restrict__ code, const float *__restrict__ table, int T,
                   const float *__restrict__ a, int lda, float *__restrict__ c, int ldc, int n, int m, float alpha, float beta) {
    asm volatile("griddepcontrol.wait;" ::: "memory");         // launched programmatically dependent on the contraction
    if (*flag != call_id) return;
    const int lane = threadIdx.x & 31;
    for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
        const int s = row_ptr[j], e = row_ptr[j + 1];
        if (e == s && beta == 1.0f) continue;
        for (int i = lane; i < m; i += 32) {
            float acc = 0.0f;
            for (int p = s; p < e; ++p) {
                const int cd = code[p];
                acc = fmaf(a[(size_t)i * lda + col[p]], cd < T ? table[cd] * alpha : 0.0f, acc);
            }
            float v = c[(size_t)i * ldc + j];
            if (beta != 1.0f) v *= beta;
            c[(size_t)i * ldc + j] = e > s ? v + acc : v;
        }
    }
}

#ifdef EXP_TC_TIMELINE
static int g_tcl_ctas = 0;
static void dump_tc_timeline() {
    static unsigned long long h[4096 * 16];
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(h, g_tc_timeline, sizeof(h));
    const int n = std::min(g_tcl_ctas, 4096);
    unsigned long long t0 = ~0ull, t7 = 0;
    for (int i = 0; i < n; ++i) { t0 = std::min(t0, h[i * 16]); t7 = std::max(t7, h[i * 16 + 7]); }
    const char *names[8] = {"start (after the first CTA)", "wait for the split kernel", "tensor memory + barrier", "first operands ready",
                            "k loop", "partial tile + counter", "fix-up / direct epilogue", "exit barrier"};
    fprintf(stderr, "[tc timeline] %d CTAs, first start to last end %.1f us\n", n, (t7 - t0) * 1e-3);
    for (int p = 0; p < 8; ++p) {
        double avg = 0, mx = 0; int cnt = 0;
        for (int i = 0; i < n; ++i) {
            const unsigned long long a = p == 0 ? t0 : h[i * 16 + p - 1], b = h[i * 16 + p];
            if (p == 5 && b == 0) continue;
            const unsigned long long a2 = (p == 6 && h[i * 16 + 5] == 0) ? h[i * 16 + 4] : a;
            if (b < a2) continue;
            const double d = (b - a2) * 1e-3; avg += d; mx = std::max(mx, d); ++cnt;
        }
        fprintf(stderr, "[tc timeline] %-30s avg %7.2f us  max %7.2f us (%d CTAs)\n", names[p], cnt ? avg / cnt : 0.0, mx, cnt);
    }
    // inside the last CTA's addition of the shares (first column group): stamps 8..12 relative to stamp 5
    const char *fx[8] = {"fence", "C loads issued", "partials arrived", "added", "stored", "prefetch + index math", "proxy fence", "copies of round 0 issued"};
    for (int p = 8; p < 16; ++p) {
        double avg = 0; int cnt = 0;
        for (int i = 0; i < n; ++i)
            if (h[i * 16 + p] > h[i * 16 + 5] && h[i * 16 + p] < h[i * 16 + 6] + 1) { avg += (h[i * 16 + p] - h[i * 16 + 5]) * 1e-3; ++cnt; }
        fprintf(stderr, "[tc timeline]   shares, first group: %-18s at %7.2f us (%d CTAs)\n", fx[p - 8], cnt ? avg / cnt : 0.0, cnt);
    }
}
#endif

template <int N_TILE, int STAGES>
int launch_tc(sblas_b200_matrix *mat, const float *xh, const float *xl, int nkb, int m, float *d_c, int ldc, float alpha, float beta,
              uint32_t *flag, uint32_t call_id, int ksplit, float *partial, uint32_t *counters, int red, cudaStream_t stream) {
    constexpr size_t smem = (size_t)STAGES * (2 * N_TILE * 128) + 256 * kTcTabCopies * 4 + (2 * STAGES + 2 * kTcWStages + 3) * 8 + 16 + 1024;
    static bool configured_on[64] = {false};
    bool &configured = configured_on[mat->device & 63];
    if (!configured) {
        SB_CUDA(cudaFuncSetAttribute(spmm_tc_kernel<N_TILE, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    const dim3 grid((mat->n + kTcM - 1) / kTcM, (m + N_TILE - 1) / N_TILE, ksplit);
    const int dbg = getenv("SBLAS_B200_TC_DBG") ? atoi(getenv("SBLAS_B200_TC_DBG")) : 0;
#ifdef EXP_TC_TIMELINE
    { static bool reg = false; if (!reg) { atexit(dump_tc_timeline); reg = true; } g_tcl_ctas = grid.x * grid.y * grid.z; }
#endif
    SB_CUDA(launch_dependent(spmm_tc_kernel<N_TILE, STAGES>, grid, dim3(kTcThreads), smem, stream, (const uint8_t *)mat->d_dense_code,
                             (int64_t)mat->dense_kpad, (const int32_t *)mat->d_row_ptr, (const float *)mat->d_table, (int)mat->T, alpha,
                             xh, xl, nkb, (int)mat->n, m, d_c, ldc, beta, (const uint32_t *)flag, call_id, ksplit, partial, counters,
                             red, dbg));
    return SBLAS_B200_OK;
}

}  // namespace

// The tile width (64 | 128 | 256) the tensor-core form would use for an m-row product of this matrix, 0 when the CUDA-core
// kernels are expected to be faster or the form does not apply.  *ksplit_out = shares of k.
static int tc_plan(const sblas_b200_matrix *mat, int m, float alpha, int *ksplit_out) {
    const char *env = getenv("SBLAS_B200_SPMM_TC");
    const int mode = (env && *env) ? atoi(env) : -1;
    if (mode == 0 || mat->dense_state < 0 || !mat->csr_sorted || !mat->table_finite || !isfinite(alpha)) return 0;
    if (mat->n < 1 || mat->k < 1 || m < 2) return 0;
    static int sms = 0;
    if (!sms && (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mat->device) != cudaSuccess || sms <= 0)) sms = 148;
    const int n_tile = m <= 64 ? 64 : (m <= 128 ? 128 : 256);
    const int mtiles = (m + n_tile - 1) / n_tile;
    const int64_t kpad = ((int64_t)mat->k + 511) / 512 * 512;   // spmv_dense.cu pads the code matrix to whole 512-column warps
    const int nkb = (int)(kpad / kTcBK);
    const int tiles = ((mat->n + kTcM - 1) / kTcM) * mtiles;
    // k is cut into shares of a quarter of the k blocks, at least 8 and at most 128 of them, each accumulated by its own CTA
    // and added in share order afterwards.  A function of k alone: a row's result then does not depend on which slab of W
    // the handle holds (nor on m within one tile width), small matrices get four times the CTAs, and no accumulator takes
    // more than 128 x 12 MMAs -- the tensor core's fp32 accumulate rounds toward zero, so its error grows with their number
    // (16384 columns in one accumulator: backward error 7e-6 measured, 2e-6 in four).
    const int share = std::min(128, std::max(8, (nkb / 4 + 3) / 4 * 4));
    int ksplit = std::max(1, (nkb + share - 1) / share);
    if (const char *e = getenv("SBLAS_B200_TC_KSPLIT")) ksplit = std::max(1, std::min(nkb / 4, atoi(e)));
    *ksplit_out = ksplit;
    if (mode != 1) {
        // Which form is faster?  Measured models, microseconds (profiles/r02_spmm_tc.md): the tile kernel moves one 128-byte
        // shared-memory wavefront per non-zero and 32 dense columns at 58 % of the pipe's rate; a tensor-core CTA spends
        // 0.42 / 0.56 / 0.92 us per 32-column step of k at N = 64 / 128 / 256, CTAs run in waves of one per SM.
        if (m < 32) return 0;
        const double t_tile = 12.0 + (double)mat->nnz * ((m + 31) / 32) / 1.69e5;
        const double t_step = n_tile == 64 ? 0.42 : (n_tile == 128 ? 0.56 : 0.92);
        // fixed part (three or four dependent launches, set-up, pipeline fill, partial tiles): 16 / 19 / 23 us measured at
        // N = 64 / 128 / 256 with the shares added by tc_fixup_kernel; 8 us more when the last CTA of a tile adds them
        const int64_t waves = ((int64_t)tiles * ksplit + sms - 1) / sms;
        const double t_tc = 14.0 + 0.035 * n_tile + (double)waves * ((nkb + ksplit - 1) / ksplit) * t_step + ((ksplit > 1 && waves > 1) ? 8.0 : 0.0);
        if (t_tc > 0.9 * t_tile) return 0;
    }
    return n_tile;
}

// 0 = done, 1 = not applicable (the caller runs the CUDA-core kernels), < 0 = error.
// SBLAS_B200_SPMM_TC = 0 never | 1 whenever possible (read per call: tests flip it).
int run_spmm_tc(sblas_b200_matrix *mat, const float *d_a, int m, int lda, float *d_c, int ldc, float alpha, float beta,
                cudaStream_t stream) {
    int ksplit = 1;
    const int n_tile = tc_plan(mat, m, alpha, &ksplit);
    if (!n_tile) return 1;
    // C in a peer GPU's memory (sblas_b200_peer_open) and beta == 1: the epilogue adds with reductions at the owner's memory
    const int red = (is_peer_pointer(d_c) && beta == 1.0f) ? 1 : 0;
    static int sms = 0;
    if (!sms && (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, mat->device) != cudaSuccess || sms <= 0)) sms = 148;
    const int mtiles = (m + n_tile - 1) / n_tile;
    const int64_t kpad = ((int64_t)mat->k + 511) / 512 * 512;
    const int nkb = (int)(kpad / kTcBK);
    const int tiles = ((mat->n + kTcM - 1) / kTcM) * mtiles;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
    const bool capturing = cap != cudaStreamCaptureStatusNone;
    if (mat->dense_state == 0) {
        if (capturing) return 1;                                // no allocation inside a capture
        const int rc = ensure_dense_codes(mat, stream);
        if (rc != SBLAS_B200_OK) return rc;                     // 1: no room for the code matrix
    }
    if (mat->dense_kpad != kpad) return fail(SBLAS_B200_ECUDA, "dense code matrix: padded width %lld, expected %lld", (long long)mat->dense_kpad, (long long)kpad);
    // workspace: [flag + pad | tile counters | X_hi | X_lo | partial tiles].  The head keeps the largest size any product of
    // this handle needed (at least 64 KB), so that the counters of a product with many tiles never lie where a product
    // with fewer tiles put its X images; they are zero between products (the last CTA of a tile puts its counter back).
    const size_t x_elems = (size_t)mtiles * nkb * n_tile * kTcBK;
    const size_t head_need = (256 + (size_t)tiles * 4 + 1023) / 1024 * 1024;
    const size_t head_pad = std::max(std::max(head_need, mat->tc_head_bytes), (size_t)65536);
    const size_t part_elems = ksplit > 1 ? (size_t)tiles * ksplit * n_tile * kTcM : 0;
    const size_t need = head_pad + (2 * x_elems + part_elems) * 4;
    if (need > mat->tc_ws_bytes || head_pad != mat->tc_head_bytes) {
        if (capturing) return 1;                                // no allocation (or re-initialisation) inside a capture
        const int rc = ensure_buffer(&mat->d_tc_ws, &mat->tc_ws_bytes, need);
        if (rc != SBLAS_B200_OK) {                              // no room for the operand images (2 x the size of A, padded, and
            cudaGetLastError();                                 // the partial tiles): the CUDA-core kernels run this product
            mat->tc_head_bytes = 0;
            return 1;
        }
        SB_CUDA(cudaMemsetAsync(mat->d_tc_ws, 0, head_pad, stream));   // flag 0 (call ids start at 1), counters 0
        mat->tc_head_bytes = head_pad;
    }
    uint8_t *ws = reinterpret_cast<uint8_t *>(mat->d_tc_ws);
    uint32_t *flag = reinterpret_cast<uint32_t *>(ws), *counters = reinterpret_cast<uint32_t *>(ws + 256);
    float *xh = reinterpret_cast<float *>(ws + head_pad), *xl = xh + x_elems, *partial = xl + x_elems;
    const uint32_t call_id = ++mat->tc_calls ? mat->tc_calls : ++mat->tc_calls;     // never 0
    const int vec = (lda % 4 == 0 && (reinterpret_cast<uintptr_t>(d_a) & 15) == 0) ? 1 : 0;
    // event pair around the product's kernels when profiling: the split kernel, the contraction and what follows are
    // chained by programmatic dependent launches and overlap, an event between them would undo that (and did: the
    // contraction alone then measured longer than the whole product)
    struct Timer {
        sblas_b200_matrix *mat; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
        Timer(sblas_b200_matrix *m_, cudaStream_t s_) : mat(m_), st(s_) {
            if (mat->profiling) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, st); }
        }
        ~Timer() { if (a) { cudaEventRecord(b, st); mat->prof_events.push_back(std::make_pair(a, b)); } }
    } timer(mat, stream);
    SB_LAUNCH(tc_split_x_kernel, dim3(nkb, mtiles, n_tile * 8 / 256), 256, 0, stream, d_a, m, mat->k, lda, n_tile, nkb, xh, xl, flag, call_id, vec);
    {
        char name[64];
        snprintf(name, sizeof(name), "spmm_tc<%d>x%d", n_tile, ksplit);
        // one code byte per ELEMENT of W is what this kernel reads
        note_product(mat, name, m, (double)mat->n * (double)mat->dense_kpad / (double)mat->nnz, (int64_t)4 * (mat->n + 1));
        int rc;
        // one wave of CTAs (small matrices, slabs of a sharded one): the shares are added by a kernel of their own
        const bool own_fixup = ksplit > 1 && (int64_t)tiles * ksplit <= sms && !(getenv("SBLAS_B200_TC_FIXUP") && atoi(getenv("SBLAS_B200_TC_FIXUP")) == 0);
        uint32_t *cnt = own_fixup ? nullptr : counters;
        if (n_tile == 64) rc = launch_tc<64, 8>(mat, xh, xl, nkb, m, d_c, ldc, alpha, beta, flag, call_id, ksplit, partial, cnt, red, stream);
        else if (n_tile == 128) rc = launch_tc<128, 6>(mat, xh, xl, nkb, m, d_c, ldc, alpha, beta, flag, call_id, ksplit, partial, cnt, red, stream);
        else rc = launch_tc<256, 3>(mat, xh, xl, nkb, m, d_c, ldc, alpha, beta, flag, call_id, ksplit, partial, cnt, red, stream);
        if (rc != SBLAS_B200_OK) return rc;
        if (own_fixup)
            SB_CUDA(launch_dependent(tc_fixup_kernel, dim3((mat->n + kTcM - 1) / kTcM, mtiles, n_tile / 8), dim3(kTcM), 0, stream, partial, ksplit,
                                     n_tile, mat->d_row_ptr, mat->n, m, d_c, ldc, beta, red, flag, call_id));
    }
    SB_CUDA(launch_dependent(tc_fallback_kernel, dim3(sms), dim3(256), 0, stream, flag, call_id, mat->d_row_ptr, mat->d_col, mat->d_code,
                             mat->d_table, mat->T, d_a, lda, d_c, ldc, mat->n, m, alpha, beta));
    SB_CHECK_LAUNCH();
    return SBLAS_B200_OK;
}

}  // namespace sblas_b200
