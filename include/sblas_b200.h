/*
 * sblas_b200.h -- C ABI of the B200 (sm_100a) implementation of the
 * NeverLEX/sparsematrix sparse x dense product path.
 *
 * This is the drop-in boundary.  The reference computes
 *      C[m x n] = beta*C + alpha * A[m x k] * S[k x n]
 * in sblas::SparseMatrix<uint8,uint8,float>::AddMatMat
 * (reference src/sparse/sparse-matrix.cc:139-194), which calls three free
 * CPU kernels with raw pointers (src/sparse/kernel.h:42-62; call sites
 * sparse-matrix.cc:150,181-189).  The entry points below are what that member
 * function binds instead: the object's compressed members go to
 * sblas_b200_upload() once, and each AddMatMat call becomes one
 * sblas_b200_addmatmat_host() (same pointers, same meaning) or, for callers
 * that keep activations on the device, sblas_b200_addmatmat_dev().
 * INTEGRATION.md shows the reference-side patch.
 *
 * Plain C: pointers and sizes only, no C++/torch types, no exceptions.
 * Every function returns an int status: SBLAS_B200_OK (0) or a negative
 * SBLAS_B200_E* code; sblas_b200_last_error() gives the thread's last message.
 * There is no CPU fallback: without a CUDA device every call that needs one
 * fails with SBLAS_B200_ECUDA.
 *
 * Thread-safety (mirrors the reference): an uploaded matrix is
 * immutable; concurrent products on one handle are serialised on the handle's
 * workspace lock; upload/free of different handles are independent.
 */
#ifndef SBLAS_B200_H_
#define SBLAS_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SBLAS_B200_API __attribute__((visibility("default")))
#else
#define SBLAS_B200_API
#endif

#define SBLAS_B200_OK        0
#define SBLAS_B200_EINVAL   -1   /* bad argument (null pointer, negative size, ld too small, ...) */
#define SBLAS_B200_ECUDA    -2   /* a CUDA call failed, or no device */
#define SBLAS_B200_ENOMEM   -3   /* host or device allocation failed */
#define SBLAS_B200_EFORMAT  -4   /* the compressed stream / CSR is inconsistent */
#define SBLAS_B200_EIO      -5   /* a file could not be opened, read or written */

#define SBLAS_B200_ABI_VERSION 1

/* Device-resident matrix W = S^T (n x k), immutable after upload.  Opaque. */
typedef struct sblas_b200_matrix sblas_b200_matrix;

/* What sblas_b200_get_info() reports. */
typedef struct sblas_b200_info {
    int32_t device;          /* CUDA ordinal the matrix lives on */
    int32_t s_rows;          /* k : rows of S (= rows_, sparse-matrix.h:50) */
    int32_t s_cols;          /* n : columns of S held by this handle (= cols_, or the slab width) */
    int32_t col_begin;       /* first S column of the slab (0 for a whole matrix) */
    int32_t table_size;      /* T : valid codebook entries */
    int32_t max_row_nnz;     /* longest row of W */
    int32_t long_rows;       /* rows handled by the split-row SpMV path */
    int32_t spmv_lanes;      /* sub-warp width the row-length histogram selected for SpMV */
    int64_t nnz;             /* stored non-zeros (fillers dropped) */
    int64_t fillers;         /* filler entries met in the stream (sparse-matrix.cc:46-51) */
    int64_t entries;         /* nnz + fillers for stream uploads, nnz for CSR uploads */
    int64_t device_bytes;    /* bytes of HBM held by the handle (format + workspace) */
    int64_t row_len_hist[33];/* [0] = empty rows, [b] = rows with 2^(b-1) <= nnz < 2^b */
} sblas_b200_info;

/* ---- library ------------------------------------------------------------ */
SBLAS_B200_API int32_t sblas_b200_abi_version(void);
SBLAS_B200_API const char *sblas_b200_last_error(void);
SBLAS_B200_API int sblas_b200_device_count(int32_t *count);
/* Kernels launched by this library in this process so far (bench.py's gpu_launches). */
SBLAS_B200_API int64_t sblas_b200_launch_count(void);

/* ---- format upload ------------------------------------------------------ */
/*
 * Upload the compressed members of a sblas::SparseMatrix<uint8,uint8,float>
 * (reference src/sparse/sparse-matrix.h:45-52, built by CopyForm,
 * sparse-matrix.cc:21-99) and convert them on the device to the layout the
 * kernels read.  Host pointers.
 *   pos_index, val_index : the two byte streams, `entries` bytes each
 *   val_table            : valid_table_size (+1 trailing zero, ignored) floats
 *   block_bounds         : nblocks pairs (row_off, col_off)  -- row_off must be 0 (block_row_shift 0)
 *   block_index_bounds   : nblocks pairs [start, end) into the streams
 *   rows, cols           : rows_ (k) and cols_ (n)
 * valid_table_size == 0 / entries == 0 give an empty matrix (sparse-matrix.cc:26).
 */
SBLAS_B200_API int sblas_b200_upload(const uint8_t *pos_index, const uint8_t *val_index, int64_t entries,
                                     const float *val_table, int32_t valid_table_size,
                                     const int32_t *block_bounds, const int32_t *block_index_bounds,
                                     int32_t nblocks, int32_t rows, int32_t cols,
                                     int32_t device, sblas_b200_matrix **out);

/*
 * Same, but keep only the S columns [col_begin, col_end) -- one row slab of W
 * for one GPU.  col_begin must be a multiple of 256 (panel aligned) so the
 * streams split without re-encoding; col_end is a multiple of 256 or == cols.
 * The product entry points then take `c` pointing at column col_begin.
 */
SBLAS_B200_API int sblas_b200_upload_slab(const uint8_t *pos_index, const uint8_t *val_index, int64_t entries,
                                          const float *val_table, int32_t valid_table_size,
                                          const int32_t *block_bounds, const int32_t *block_index_bounds,
                                          int32_t nblocks, int32_t rows, int32_t cols,
                                          int32_t col_begin, int32_t col_end,
                                          int32_t device, sblas_b200_matrix **out);

/*
 * The same for sblas::SparseMatrix<uint16, uint8, float>: position deltas of TWO bytes
 * (PosIndex_t = uint16, zero_pad_interval 65535, sparse-matrix.cc:24), so hypersparse
 * matrices need no fillers below gaps of 65 535 (0.1 % density: 2 bytes + 1 per non-zero
 * instead of 8.9).  The reference has the code path but never instantiates it
 * (sparse-matrix.cc:315).  A whole matrix is col_begin = 0,
 * col_end = cols.  Everything after the upload (device layouts, kernels) is the same.
 */
SBLAS_B200_API int sblas_b200_upload_slab16(const uint16_t *pos_index, const uint8_t *val_index, int64_t entries,
                                            const float *val_table, int32_t valid_table_size,
                                            const int32_t *block_bounds, const int32_t *block_index_bounds,
                                            int32_t nblocks, int32_t rows, int32_t cols,
                                            int32_t col_begin, int32_t col_end,
                                            int32_t device, sblas_b200_matrix **out);

/*
 * Direct CSR upload of W (n x k) for shapes the panel format cannot hold
 * (e.g. 2^23 rows: SURVEY.md 0.6c).  row_ptr[n+1], col[nnz] ascending inside a
 * row, code[nnz] < T indexing table[T].  Host pointers.
 */
SBLAS_B200_API int sblas_b200_upload_csr(const int32_t *row_ptr, const int32_t *col, const uint8_t *code,
                                         const float *table, int32_t valid_table_size,
                                         int32_t n, int32_t k, int32_t device, sblas_b200_matrix **out);

/*
 * nnz-balanced, panel-aligned split of the S columns into `parts` contiguous
 * slabs: col_splits[parts+1], col_splits[0] = 0,
 * col_splits[parts] = cols, interior values multiples of 256.  Pure host code.
 */
SBLAS_B200_API int sblas_b200_partition(const int32_t *block_bounds, const int32_t *block_index_bounds,
                                        int32_t nblocks, int32_t cols, int32_t parts, int32_t *col_splits);

/*
 * The same for a matrix given as CSR of W (sblas_b200_upload_csr): row_splits[parts+1],
 * nnz-balanced, interior values multiples of `align` (a power of two; 256 keeps the slabs
 * panel aligned like the stream form).  Slab g is rows [row_splits[g], row_splits[g+1]):
 * upload it with row_ptr + row_splits[g] rebased to 0.  Pure host code.
 */
SBLAS_B200_API int sblas_b200_partition_csr(const int32_t *row_ptr, int32_t n, int32_t parts, int32_t align,
                                            int32_t *row_splits);

/*
 * CopyForm on the device (reference src/sparse/sparse-matrix.cc:21-99): dense
 * code matrix -> the reference's byte streams, bit for bit (fillers, per-panel
 * restart of the position, empty panels not recorded), and from them the device
 * layout without a host round trip.  Two calls because the caller owns the
 * output arrays and their sizes are a result:
 *   _begin  : codes is rows x stride bytes (host or device pointer), a code
 *             < valid_table_size is an element, anything else is a zero; trans != 0
 *             means the input is S^T (SblasTrans).  Returns the job, the number
 *             of stream entries and of recorded panels.
 *   _finish : writes pos_index/val_index (entries bytes each), block_bounds and
 *             block_index_bounds (nblocks pairs each) to HOST arrays; when `mat`
 *             is not NULL also builds the device matrix (val_table: valid_table_size
 *             floats) as sblas_b200_upload would.  Frees the job in every case.
 *   _abort  : frees a job that will not be finished.
 */
typedef struct sblas_b200_copyform_job sblas_b200_copyform_job;
SBLAS_B200_API int sblas_b200_copyform_begin(const uint8_t *codes, int32_t rows, int32_t cols, int32_t stride,
                                             int32_t valid_table_size, int32_t trans, int32_t device,
                                             sblas_b200_copyform_job **job, int64_t *entries, int32_t *nblocks);
SBLAS_B200_API int sblas_b200_copyform_finish(sblas_b200_copyform_job *job, uint8_t *pos_index, uint8_t *val_index,
                                              int32_t *block_bounds, int32_t *block_index_bounds,
                                              const float *val_table, sblas_b200_matrix **mat);
SBLAS_B200_API void sblas_b200_copyform_abort(sblas_b200_copyform_job *job);

SBLAS_B200_API int sblas_b200_free(sblas_b200_matrix *mat);
SBLAS_B200_API int sblas_b200_get_info(const sblas_b200_matrix *mat, sblas_b200_info *info);

/* ---- the product (replaces the body of AddMatMat, sparse-matrix.cc:139-194) ---- */
/*
 * C = beta*C + alpha*A*S with HOST pointers, the reference's own signature:
 * a is m x k row-major with leading dimension lda, c is m x n row-major with
 * leading dimension ldc, updated in place.  Copies a and c to the device, runs
 * the kernels, copies c back, and returns when c is valid.  Semantics kept
 * from the reference: beta == 1 skips the scale, beta == 0 still multiplies
 * (NaN/Inf in C propagate), alpha == 0 only scales, S columns that hold no
 * entry only get the scale.
 */
SBLAS_B200_API int sblas_b200_addmatmat_host(sblas_b200_matrix *mat, const float *a, int32_t m, int32_t lda,
                                             float *c, int32_t ldc, float alpha, float beta);

/*
 * The same product on DEVICE pointers (on mat's device), enqueued on `stream`
 * (a cudaStream_t; NULL = the default stream) and not synchronised.  This is
 * the timed path.  Calls on one handle may come from several threads and
 * streams: the handle's workspace is shared, so a product enqueued on another
 * stream than the previous one is ordered after it with an event (products on
 * different handles run concurrently).  The call can be recorded into a CUDA
 * graph (stream capture) once each shape has been run eagerly -- workspaces are
 * then allocated and no allocation or synchronisation happens inside; replays of
 * such a graph and other products on the same handle must be ordered by the
 * caller, the library's event cannot cross a capture.  With pinned host buffers
 * sblas_b200_addmatmat_host() overlaps its copies with the product.
 */
SBLAS_B200_API int sblas_b200_addmatmat_dev(sblas_b200_matrix *mat, const float *d_a, int32_t m, int32_t lda,
                                            float *d_c, int32_t ldc, float alpha, float beta, void *stream);

/*
 * The same product for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  This entry
 * point has NO counterpart in the reference.
 * It calls no cudaMalloc / cudaFree / allocating routine, ever:
 * the product runs as transpose (X = A^T) + the gather-form kernel on the CSR the
 * upload left on the device, with X staged in ONE fixed device array of
 * SBLAS_B200_NOALLOC_MAX_FLOATS floats (4 MiB) that every GPU which loads the
 * library holds until exit.  Limit: k * round_up(m, 4) <= SBLAS_B200_NOALLOC_MAX_FLOATS
 * (e.g. k = 4096 with m <= 256, k = 16384 with m <= 64); beyond it the call fails
 * with SBLAS_B200_EINVAL and a message naming the limit, and nothing is enqueued.
 * All calls on one device share the array: calls that overlap on the GPU
 * (different streams, or different handles, without ordering between them) RACE on it
 * -- keep them on one stream or order them yourself.  The array keeps the previous
 * call's contents; nothing in it needs to start at zero, each call rewrites every
 * float it later reads (padding columns included) before reading it.  Arguments,
 * semantics and results are those of sblas_b200_addmatmat_dev() (results may differ
 * from it in rounding where that picks another kernel: same 1e-5 backward bound).
 */
#define SBLAS_B200_NOALLOC_MAX_FLOATS (1 << 20)
SBLAS_B200_API int sblas_b200_addmatmat_dev_noalloc(sblas_b200_matrix *mat, const float *d_a, int32_t m, int32_t lda,
                                                    float *d_c, int32_t ldc, float alpha, float beta, void *stream);

/* ---- the reference's free kernel functions (src/sparse/kernel.h:42-62) ------ */
/*
 * For callers that use the reference's free functions directly.  Host pointers,
 * synchronous, on the current device; sparsematrix_b200/src/sparse/kernel.{h,cc}
 * bind the reference's own template names to them.
 *   _scale_host         : c[m x n, ldc] *= beta                 (sblas_beta_operation_kernel, kernel.cc:10-29)
 *   _transpose_host     : sa[j*ldsa + i] = a[i*lda + j], i < m, j < n; nothing else of sa is written
 *                                                               (sblas_trans_kernel, kernel.cc:31-187)
 *   _panel_product_host : C += alpha * A * P for ONE panel P (k x n, n <= 256) given as its byte streams
 *                         ppos/pval of pos_len entries.  trans == 0: a is m x k, c is m x n, row-major
 *                         (sblas_kernel_operation, _naive, kernel.cc:213-369); trans != 0: a is k x lda and
 *                         c is n x ldc with the m values contiguous (sblas_kernel_operation_trans, _trans_ex,
 *                         kernel.cc:612-800, the layout AddMatMat hands them, sparse-matrix.cc:181-189).
 *                         Uploads the panel for the call: a convenience, not the fast path.
 */
SBLAS_B200_API int sblas_b200_scale_host(float *c, int32_t m, int32_t n, int32_t ldc, float beta);
SBLAS_B200_API int sblas_b200_transpose_host(const float *a, int32_t m, int32_t n, int32_t lda, float *sa, int32_t ldsa);
SBLAS_B200_API int sblas_b200_panel_product_host(int32_t trans, int32_t m, int32_t n, int32_t k, const float *a, int32_t lda,
                                                 float *c, int32_t ldc, float alpha, const uint8_t *ppos, const uint8_t *pval,
                                                 int32_t pos_len, const float *val_table, int32_t valid_table_size);

/*
 * Page-locked, device-mapped host memory for the caller's a and c.  Any host
 * pointer works with sblas_b200_addmatmat_host(); with memory of this kind
 * (these calls, cudaHostAlloc or cudaHostRegister) the copies run beside the
 * product and the result is written straight into c.  Fails with
 * SBLAS_B200_ECUDA without a device: there is no malloc fallback.
 */
SBLAS_B200_API int sblas_b200_host_alloc(int64_t bytes, void **ptr);
/*
 * Pageable (malloc) buffers: what an unmodified caller of the reference passes
 * (src/test/blas_test.h:121).  They are staged through the driver (cfg2: 570 us per
 * call against 157 with the memory above).  With SBLAS_B200_HOST_REGISTER=1 in the
 * environment (read at every call) a buffer that comes back a SECOND time at the same
 * address with the same size (>= 64 KiB) is page-locked once with cudaHostRegister
 * and then takes the overlapped route (cfg2: 179 us).  The registration belongs to
 * the handle of that call: sblas_b200_free() drops it, at most 4 GiB stay registered
 * (oldest out first).  Opt-in because of cudaHostRegister's own CONTRACT: a registered
 * buffer must not be unmapped (free() of a large block) unless
 * sblas_b200_host_forget(ptr) (NULL = every range) is called first -- a new allocation
 * at the same address would otherwise be read through the old pages.  For callers that
 * keep their buffers.  No counterpart in the reference.
 */
SBLAS_B200_API int sblas_b200_host_forget(const void *ptr);
SBLAS_B200_API int sblas_b200_host_free(void *ptr);

/* ---- peer-visible output: reassembly across GPUs without a collective ------ */
/*
 * The slabs of sblas_b200_upload_slab() write disjoint column ranges of C (the
 * reference's panel loop, sparse-matrix.cc:164-190, has the same independence),
 * so the whole C can be assembled on one GPU by letting every rank's product
 * kernels store through an NVLink-mapped pointer: d_c = base + col_begin,
 * ldc = n in sblas_b200_addmatmat_dev().  One process per GPU, so the mapping
 * goes through a CUDA IPC handle (64 opaque bytes, sent by any channel):
 *   _peer_alloc : cudaMalloc on `device` + the handle of that allocation
 *   _peer_open  : map another process's allocation of `bytes` bytes into this process
 *                 for `device` (peer access is enabled on first use); not for the
 *                 owner itself.  Products whose c lies inside such a mapping skip
 *                 the L2 prefetch of C and, for beta == 1, add their sums with
 *                 reductions at the owner's memory instead of reading C over the link
 *                 (same single rounding; subnormal sums are flushed to zero there)
 *   _peer_close : unmap a pointer of _peer_open
 *   _peer_free  : free a pointer of _peer_alloc (after every peer has closed it)
 *   _peer_register / _peer_unregister : declare a range of peer memory the application
 *                 mapped by other means (one process driving several GPUs with
 *                 cudaDeviceEnablePeerAccess, cuMemMap of a fabric handle, a symmetric
 *                 heap), so that products into it take the same epilogues.  Nothing is
 *                 mapped or freed; ranges must not overlap.  No device is needed.
 * The owner must not read C before the writers' streams have been synchronised
 * (an application-level barrier); stores to peer memory are not ordered with
 * the owner's own stream.
 */
#define SBLAS_B200_IPC_HANDLE_BYTES 64
SBLAS_B200_API int sblas_b200_peer_alloc(int32_t device, int64_t bytes, void **d_ptr, uint8_t *handle);
SBLAS_B200_API int sblas_b200_peer_open(int32_t device, const uint8_t *handle, int64_t bytes, void **d_ptr);
SBLAS_B200_API int sblas_b200_peer_close(void *d_ptr);
SBLAS_B200_API int sblas_b200_peer_free(void *d_ptr);
SBLAS_B200_API int sblas_b200_peer_register(void *d_ptr, int64_t bytes);
SBLAS_B200_API int sblas_b200_peer_unregister(void *d_ptr);

/* ---- kernel timing (bench.py's roofline leg) ------------------------------ */
/*
 * Between _begin and _end every product on this handle brackets its dominant
 * kernel (the SpMV row kernel for m == 1, the SpMM kernel otherwise -- not the
 * transpose or the split-row fix-ups) with a CUDA event pair on the launching
 * stream.  _end synchronises those events and returns the summed kernel time
 * and the number of bracketed launches.
 */
SBLAS_B200_API int sblas_b200_profile_begin(sblas_b200_matrix *mat);
SBLAS_B200_API int sblas_b200_profile_end(sblas_b200_matrix *mat, double *total_ms, int64_t *launches);

/*
 * What the most recent product on this handle ran: the name of its dominant
 * kernel (the one _profile_begin/_end bracket) and the bytes that kernel has to
 * move once per launch for the device format it reads -- the "algorithmic bytes"
 * of the roofline: format bytes per stored non-zero, row pointers,
 * X once, C read and written once.  m is the dense row count of that product.
 */
typedef struct sblas_b200_product_info {
    char kernel[64];
    int32_t m;
    int32_t reserved;
    int64_t algorithmic_bytes;
    int64_t algorithmic_flops;    /* 2 * nnz * m */
    double format_bytes_per_nnz;  /* 5 for CSR with codes and the quad-row tiles, 13 for the two-pass blocked SpMV, ... */
} sblas_b200_product_info;
SBLAS_B200_API int sblas_b200_last_product(const sblas_b200_matrix *mat, sblas_b200_product_info *out);

/* fp32 FMA throughput of `device` measured with a register-resident FMA loop (TFLOP/s): the roofline's compute peak. */
SBLAS_B200_API int sblas_b200_measure_fp32_peak(int32_t device, double *tflops);

/* ---- structure read-back (parity tests) --------------------------------- */
/*
 * Decode on the device: the function of CopyTo (sparse-matrix.cc:101-137) for
 * the columns this handle holds.  `dense` is a host pointer (decoded into a
 * staging buffer and copied out) or a device pointer (decoded in place).
 *   trans == 0 : dense is s_rows x stride, dense[r*stride + j] = S(r, col_begin + j)
 *   trans == 1 : dense is s_cols x stride, dense[j*stride + r] = S(r, col_begin + j)
 * The whole leading x stride area is cleared first, like the reference's memset.
 */
SBLAS_B200_API int sblas_b200_decode_dense(const sblas_b200_matrix *mat, float *dense, int32_t stride, int32_t trans);

/* Copy the device CSR of W back: row_ptr[s_cols+1], col[nnz], code[nnz] (host pointers). */
SBLAS_B200_API int sblas_b200_export_csr(const sblas_b200_matrix *mat, int32_t *row_ptr, int32_t *col, uint8_t *code);

#ifdef __cplusplus
}
#endif
#endif /* SBLAS_B200_H_ */
