// device_matrix.h -- internal (C++/CUDA) view of the opaque sblas_b200_matrix handle,
// error plumbing and launch accounting shared by src/utils (upload, allocation)
// and src/sparse (product kernels).  Not installed; the public surface is
// include/sblas_b200.h.
#ifndef SBLAS_B200_DEVICE_MATRIX_H_
#define SBLAS_B200_DEVICE_MATRIX_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <mutex>
#include <utility>
#include <vector>

#include "sblas_b200.h"

namespace sblas_b200 {

// Panel geometry of the reference format (reference src/sparse/kernel.h:25-26).
constexpr int kPanelShift = 8;
constexpr int kPanelWidth = 1 << kPanelShift;
constexpr int kPanelMask = kPanelWidth - 1;

// SpMV plan constants (DESIGN.md "K-spmv").
constexpr int kLongSegment = 8192;   // a long row is cut into segments of this many non-zeros, one CTA each

void set_error(const char *fmt, ...);
int fail(int code, const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);

extern std::atomic<int64_t> g_kernel_launches;

// RAII: make `device` current for the duration of an entry point, restore on exit.
class DeviceScope {
public:
    explicit DeviceScope(int device);
    ~DeviceScope();
    bool ok() const { return ok_; }
private:
    int prev_;
    bool ok_;
};

}  // namespace sblas_b200

#define SB_CUDA(call)                                                                      \
    do {                                                                                   \
        cudaError_t sb_e_ = (call);                                                        \
        if (sb_e_ != cudaSuccess) return ::sblas_b200::cuda_fail(sb_e_, #call, __FILE__, __LINE__); \
    } while (0)

// Launch + account.  Every kernel this library runs goes through here, so
// sblas_b200_launch_count() is the number of our kernels that were enqueued.
#define SB_LAUNCH(kernel, grid, block, smem, stream, ...)                                  \
    do {                                                                                   \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                        \
        ::sblas_b200::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);           \
    } while (0)

#define SB_CHECK_LAUNCH() SB_CUDA(cudaGetLastError())

// One SpMM tile layout: row groups of `rows` W rows x chunks of 255 X rows; per (group, chunk) a
// contiguous run of 80-byte quad rows (per warp slot 4 fp32 weights + 4 one-byte X-row offsets).
struct TileFormat {
    int state = 0;                      // 0 = not built yet, 1 = built, -1 = not worthwhile / not possible
    int rows = 0;                       // W rows per group: 4 or 8
    uint8_t *d_blocks = nullptr;        // [nblocks + 2][80] quad rows
    int32_t *d_grp_ptr = nullptr;       // [ngroups * nchunks + 1] first quad row of each (group, chunk)
    uint32_t *d_grp_nq = nullptr;       // [ngroups * nchunks] quad rows of slot-row 0 | slot-row 1 << 16
    int32_t ngroups = 0, nchunks = 0;
    int64_t nblocks = 0;
};

// Column-blocked SpMV layout for matrices whose x gathers would each cost an L2 sector (k beyond any L1, columns
// spread over all of it): the entries of the rows up to long_threshold, sorted by (column block, row, column).
// Pass 1 streams them with the block's x tile in shared memory and writes one product per entry, pass 2 adds
// the products of a row sub-block (one warp) in stream order (src/utils/blocked_format.cu, src/sparse/spmv_blocked.cu).
struct BlockedFormat {
    int state = 0;                      // 0 = not built, 1 = built, -1 = not applicable to this matrix
    int cbw = 0;                        // columns per block = floats of x held in shared memory by pass 1
    int ncb = 0;                        // column blocks
    int rs = 0, nsub = 0;               // rows per sub-block (one warp of pass 2), number of sub-blocks
    int64_t total = 0;                  // stream slots: a column block starts on a multiple of 32, a sub-segment on one of 4
    int64_t entries = 0;                // real entries among them
    int64_t runs = 0;                   // maximal groups of stream neighbours that share a row
    uint16_t *d_lcol = nullptr;         // [total] column inside the block
    uint16_t *d_lrow = nullptr;         // [total] row inside the sub-block (0xffff in padding slots)
    uint8_t *d_code = nullptr;          // [total] codebook index (255 in padding slots)
    int32_t *d_seg_ptr = nullptr;       // [ncb][nsub + 1] first slot of (block, sub-block); [.][nsub] = end of the block's entries
    int32_t *d_cb_start = nullptr;      // [ncb + 1] first slot of a block
    float *d_prod = nullptr;            // [total] products, written by pass 1 and read by pass 2 of the same call
};

// The reference's own streams kept on the device ("native-stream kernels"): pos_index_ / val_index_ of the
// slab, byte for byte as CopyForm wrote them (sparse-matrix.cc:21-99), plus what the upload computes anyway -- segments of
// 2048 stream entries that never cross a panel, and the linear position each segment starts from (the running sum of the
// deltas before it).  src/sparse/spmv_native.cu multiplies straight from them: 2 bytes per stream entry.
struct NativeFormat {
    int state = 0;                      // 0 = not kept, 1 = kept
    uint8_t *d_pos = nullptr;           // [entries + pad] one-byte position deltas
    uint8_t *d_code = nullptr;          // [entries + pad] codebook indices (>= T: filler)
    int32_t *d_seg_begin = nullptr;     // [nseg] first / one past the last stream entry of a segment
    int32_t *d_seg_end = nullptr;
    uint32_t *d_seg_pos0 = nullptr;     // [nseg] linear position (row * 256 + column in panel) before the segment's first entry
    int32_t *d_panel_seg_first = nullptr;   // [npanels + 1] first segment of a recorded panel
    int32_t *d_slot_panel = nullptr;    // [nslots] recorded panel of the 256 columns starting at 256 * slot, or -1
    float *d_partial = nullptr;         // [nseg][256] column sums of a segment, written and read by one product
    int nseg = 0, npanels = 0, nslots = 0;
    int max_col_count = 0;              // most entries of one column inside one segment
};

struct sblas_b200_matrix {
    int device = 0;
    int32_t k = 0;           // rows of S  = length of an X row index space
    int32_t n = 0;           // columns of S held here = rows of W
    int32_t col_begin = 0;   // first S column of the slab
    int32_t T = 0;           // valid codebook entries
    int64_t nnz = 0, fillers = 0, entries = 0;

    // CSR of W = S^T, codebook-indexed values: 4 B column + 1 B code per non-zero.
    // col/code are padded to a multiple of 4 entries (+4) so rows can be read as
    // aligned 128-bit / 32-bit blocks.
    int32_t *d_row_ptr = nullptr;   // [n+1]
    int32_t *d_col = nullptr;       // [nnz_padded]
    uint8_t *d_code = nullptr;      // [nnz_padded]
    float *d_table = nullptr;       // [256], entries >= T are 0

    // SpMV plan, chosen from the row-length histogram at upload.
    int spmv_lanes = 8;             // sub-warp width for the row-split kernel
    int32_t long_threshold = INT32_MAX;  // rows longer than this take the split-row path
    int32_t n_long = 0, n_long_seg = 0, max_row_nnz = 0;
    // Skewed row lengths only: a second copy of the short rows, sorted by decreasing length, every
    // row starting on a block of 4 entries (tail padded with col 0 / code 255 = weight 0, never used).
    int32_t *d_row_order = nullptr;     // [n_ordered]     original row of sorted position i
    int32_t *d_sblk_ptr = nullptr;      // [n_ordered + 1] first block (4 entries) of sorted position i
    int32_t *d_slen = nullptr;          // [n_ordered]     entries of that row
    int4 *d_scol = nullptr;             // [sorted_blocks] column indices
    uint32_t *d_scode = nullptr;        // [sorted_blocks] four codes
    int32_t n_ordered = 0;
    int64_t sorted_blocks = 0;
    int32_t *d_long_rows = nullptr;     // [n_long]      row index
    int32_t *d_long_seg_ptr = nullptr;  // [n_long+1]    first segment of each long row
    int32_t *d_seg_begin = nullptr;     // [n_long_seg]  element range of a segment
    int32_t *d_seg_end = nullptr;       // [n_long_seg]
    int64_t row_len_hist[33] = {0};

    // Per-handle workspace (X = A^T staging, split-row partials, host-path staging).
    std::mutex lock;
    float *d_ws = nullptr;      size_t ws_bytes = 0;       // X^T [k][m_pad]
    float *d_partial = nullptr; size_t partial_bytes = 0;  // [n_long_seg * m]
    float *d_host_a = nullptr;  size_t host_a_bytes = 0;   // host path: A on device
    float *d_host_c = nullptr;  size_t host_c_bytes = 0;   // host path: C on device
    cudaStream_t stream = nullptr;                          // host path stream
    // the workspace is shared by every product of the handle: a product enqueued on another stream than the
    // previous one first waits for that one's event (same stream: stream order is enough)
    cudaEvent_t ws_event = nullptr;
    cudaStream_t ws_stream = nullptr;
    bool ws_used = false;
    // host path, pinned caller buffers: the product runs into d_host_p (kept at -0.0f between calls)
    // while C travels in on stream2; one kernel then combines both straight into the caller's C
    float *d_host_p = nullptr;  size_t host_p_bytes = 0;   bool host_p_clean = false;
    // SpMV of the length-sorted copy into peer memory: row sums formed here (kept at -0.0f), added out in row order
    float *d_peer_y = nullptr;  size_t peer_y_bytes = 0;   bool peer_y_clean = false;
    // and, for large m, in slices of A/C rows so that copies in, products and copies out overlap
    cudaStream_t stream2 = nullptr, stream3 = nullptr;  // copies in / copies out
    // SpMV with long rows: their segment sums are formed on a side stream beside the kernel of the other rows
    cudaStream_t side_stream = nullptr;
    cudaEvent_t side_fork = nullptr, side_join = nullptr;
    std::vector<cudaEvent_t> host_events;                // per slice: A arrived, C arrived, result ready
    int64_t format_bytes = 0;

    // SpMM tile formats (src/utils/tile_format.cu), built on the first large-m product that wants
    // them: [0] row groups of 4 W rows (one per warp slot), [1] groups of 8 (two per slot).
    bool csr_sorted = true;             // columns ascending inside every row (required by the tile build)
    TileFormat tile[2];

    // Dense code matrix of W (src/sparse/spmv_dense.cu), built on the first m = 1 product of a matrix
    // with >= 3/16 of its elements stored: n x dense_kpad one-byte codes, >= T where there is no entry.
    uint8_t *d_dense_code = nullptr;
    int64_t dense_kpad = 0;
    int dense_state = 0;                // 0 = not built yet, 1 = built, -1 = not possible (memory)

    BlockedFormat blocked;
    NativeFormat native;

    // Tensor-core SpMM (src/sparse/spmm_tc.cu): workspace [flag | tile counters | X_hi | X_lo | split-k partial tiles]
    bool table_finite = true;           // every codebook entry is finite (else the matrix stays off that path)
    float *d_tc_ws = nullptr;   size_t tc_ws_bytes = 0, tc_head_bytes = 0;
    uint32_t tc_calls = 0;              // id of the latest product on that path (its non-finite-x flag value)

    // sblas_b200_last_product: what the most recent product launched as its dominant kernel
    char last_kernel[64] = "";
    int32_t last_m = 0;
    int64_t last_bytes = 0;
    double last_bytes_per_nnz = 0.0;

    // sblas_b200_profile_begin/_end: event pairs around the dominant kernel
    bool profiling = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t> > prof_events;
};

namespace sblas_b200 {
// src/utils/device_format.cu
int plan_spmv(sblas_b200_matrix *mat, const int32_t *h_row_ptr);
int ensure_buffer(float **buf, size_t *have, size_t need);
int exclusive_scan_i32(const int32_t *d_in, int n, int32_t *d_out, cudaStream_t stream);   // d_out[n] = total
// src/utils/peer_memory.cu: does p point into a mapping made by sblas_b200_peer_open (another GPU's memory)?
bool is_peer_pointer(const void *p);
// src/utils/tile_format.cu
constexpr int kTileK = 255;          // X rows per chunk; offset 255 addresses an all-zero row (padding entries)
constexpr int kBlockBytes = 80;
constexpr int kTilePad = 48;         // quad rows of slack after the last one (ring segments fetched ahead)
int build_tile_format(sblas_b200_matrix *mat, int rows_per_group, cudaStream_t stream);
// src/sparse/spmm_tile.cu
int run_spmm_tile(sblas_b200_matrix *mat, const float *d_a, int m, int lda, float *d_c, int ldc, float alpha,
                  float beta, cudaStream_t stream);
// src/sparse/spmv_dense.cu: 0 = done, 1 = not applicable (use the row kernels), < 0 = error
int run_spmv_dense(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta, cudaStream_t stream);
// builds the one-byte dense code matrix of W when it does not exist yet: 0 = there, 1 = no room, < 0 = error
int ensure_dense_codes(sblas_b200_matrix *mat, cudaStream_t stream);
// src/sparse/spmm_tc.cu: 0 = done, 1 = not applicable (use the CUDA-core kernels), < 0 = error
int run_spmm_tc(sblas_b200_matrix *mat, const float *d_a, int m, int lda, float *d_c, int ldc, float alpha, float beta,
                cudaStream_t stream);
constexpr int kBlockedMaxRows = 1792;   // rows per sub-block at most: 16 warps x 1792 floats = 112 KB, two CTAs per SM
// src/utils/blocked_format.cu: 0 = built or not applicable (see blocked.state), < 0 = error
int build_blocked_format(sblas_b200_matrix *mat);
// src/sparse/spmv_blocked.cu: the rows up to long_threshold; the caller adds the long rows afterwards (beta already applied)
int run_spmv_blocked(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta, bool red, cudaStream_t stream);
// src/sparse/spmv_native.cu: 0 = done, 1 = not applicable / not chosen (use the other kernels), < 0 = error
int run_spmv_native(sblas_b200_matrix *mat, const float *x, float *y, float alpha, float beta, cudaStream_t stream);
// records the dominant kernel of the running product: name, bytes of the device format per non-zero, and the bytes per
// launch that do not scale with nnz besides X and C (row pointers, segment pointers, ...)
void note_product(sblas_b200_matrix *mat, const char *kernel, int m, double bytes_per_nnz, int64_t index_bytes);
// src/utils/pinned_memory.cu: registration cache of pageable caller buffers (second sighting of (address, size) -> cudaHostRegister)
void note_host_buffer(const void *owner, const void *ptr, size_t bytes);
void forget_host_buffers(const void *owner, const void *ptr);      // ptr != null: that range; else everything `owner` registered (null: all)
// src/sparse/product.cu
int product_on_stream(sblas_b200_matrix *mat, const float *d_a, int32_t m, int32_t lda, float *d_c, int32_t ldc,
                      float alpha, float beta, cudaStream_t stream);
// Programmatic dependent launch: the kernel may be set up while the one before it in the stream drains; it reads what
// that kernel wrote only after its own griddepcontrol.wait.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_dependent(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
    g_kernel_launches.fetch_add(1, std::memory_order_relaxed);
    return e;
}


}  // namespace sblas_b200

#endif
