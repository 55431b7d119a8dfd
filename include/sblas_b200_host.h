/*
 * sblas_b200_host.h -- C handle API over the host-side class sblas::SparseMatrix
 * <uint8,uint8,float> of this build (sparsematrix_b200/src/sparse/sparse-matrix.h,
 * the mirror of reference src/sparse/sparse-matrix.h:25-53).
 *
 * For language bindings (ctypes, cgo, JNI ...) that want the reference's object
 * interface -- CopyForm / CopyTo / AddMatMat / NumRows / NumCols / Destroy -- rather
 * than the raw device entry points of sblas_b200.h.  One function per member
 * function, same argument order and meaning.  AddMatMat runs on the GPU.
 */
#ifndef SBLAS_B200_HOST_H_
#define SBLAS_B200_HOST_H_

#include <stdint.h>

#include "sblas_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sblas_host_matrix sblas_host_matrix;

SBLAS_B200_API sblas_host_matrix *sblas_host_create(void);                 /* SparseMatrix()            sparse-matrix.h:28 */
SBLAS_B200_API void sblas_host_free(sblas_host_matrix *h);                 /* ~SparseMatrix()           :32 */
SBLAS_B200_API void sblas_host_destroy(sblas_host_matrix *h);              /* Destroy()                 :34 */
/* CopyForm(density_matrix, rows, cols, stride, vals, val_table_size, trans)  :35 ; trans: 0 NoTrans, 1 Trans */
SBLAS_B200_API int sblas_host_copyform(sblas_host_matrix *h, const uint8_t *density_matrix, int32_t rows, int32_t cols,
                                       int32_t stride, const float *vals, int32_t val_table_size, int32_t trans);
/* CopyTo(density_matrix, stride, trans)                                      :36 */
SBLAS_B200_API void sblas_host_copyto(sblas_host_matrix *h, float *density_matrix, int32_t stride, int32_t trans);
/* AddMatMat(a, m, lda, c, ldc, alpha, beta)                                  :37 ; returns the C-ABI status */
SBLAS_B200_API int sblas_host_addmatmat(sblas_host_matrix *h, const float *a, int32_t m, int32_t lda, float *c,
                                        int32_t ldc, float alpha, float beta);
SBLAS_B200_API int32_t sblas_host_num_rows(const sblas_host_matrix *h);    /* NumRows()                 :39 */
SBLAS_B200_API int32_t sblas_host_num_cols(const sblas_host_matrix *h);    /* NumCols()                 :40 */
SBLAS_B200_API int32_t sblas_host_equal(sblas_host_matrix *a, sblas_host_matrix *b); /* operator==      :42 */
SBLAS_B200_API int32_t sblas_host_selftest(void);                          /* SelfTest()                :43 (needs a GPU) */

/* The compressed members (sparse-matrix.h:45-52), for uploads, slabs and structure tests. */
SBLAS_B200_API int64_t sblas_host_entries(const sblas_host_matrix *h);
SBLAS_B200_API int32_t sblas_host_nblocks(const sblas_host_matrix *h);
SBLAS_B200_API int32_t sblas_host_table_size(const sblas_host_matrix *h);  /* valid entries T (val_table_ holds T+1) */
/* Borrowed pointers, valid until the next CopyForm / Destroy / free.  Any may be NULL to skip. */
SBLAS_B200_API void sblas_host_members(const sblas_host_matrix *h, const uint8_t **pos_index, const uint8_t **val_index,
                                       const float **val_table, const int32_t **block_bounds,
                                       const int32_t **block_index_bounds);
/*
 * Dump / load of the compressed members (the reference has no on-disk form), so that
 * weights and test fixtures travel between boxes without repeating CopyForm.  Little-endian file:
 *   "SBLASM1\0" | int32 rows, cols, valid_table_size, nblocks | int64 entries |
 *   val_table[T+1] f32 | block_bounds[nblocks][2] i32 | block_index_bounds[nblocks][2] i32 |
 *   pos_index[entries] u8 | val_index[entries] u8 | uint64 FNV-1a of everything before it.
 * _load validates sizes, bounds, codes and the checksum before touching the object (SBLAS_B200_EFORMAT
 * otherwise, SBLAS_B200_EIO when the file cannot be read), then replaces its contents like CopyForm does.
 */
SBLAS_B200_API int sblas_host_save(const sblas_host_matrix *h, const char *path);
SBLAS_B200_API int sblas_host_load(sblas_host_matrix *h, const char *path);
/* The device matrix of this object on the current CUDA device (uploads on first use); NULL on failure. */
SBLAS_B200_API sblas_b200_matrix *sblas_host_device_matrix(sblas_host_matrix *h);

/*
 * The same object interface for sblas::SparseMatrix<uint16, uint8, float> -- two-byte position deltas (PosIndex_t = uint16,
 * reference sparse-matrix.cc:24: zero_pad_interval 65535; the reference never instantiates it, :315).
 * Hypersparse matrices need no fillers below gaps of 65 535.  CopyForm runs on the host for this instantiation; AddMatMat
 * uploads through sblas_b200_upload_slab16 and runs the same kernels.
 */
typedef struct sblas_host_matrix16 sblas_host_matrix16;
SBLAS_B200_API sblas_host_matrix16 *sblas_host16_create(void);
SBLAS_B200_API void sblas_host16_free(sblas_host_matrix16 *h);
SBLAS_B200_API int sblas_host16_copyform(sblas_host_matrix16 *h, const uint8_t *density_matrix, int32_t rows, int32_t cols,
                                         int32_t stride, const float *vals, int32_t val_table_size, int32_t trans);
SBLAS_B200_API void sblas_host16_copyto(sblas_host_matrix16 *h, float *density_matrix, int32_t stride, int32_t trans);
SBLAS_B200_API int sblas_host16_addmatmat(sblas_host_matrix16 *h, const float *a, int32_t m, int32_t lda, float *c, int32_t ldc,
                                          float alpha, float beta);
SBLAS_B200_API int32_t sblas_host16_num_rows(const sblas_host_matrix16 *h);
SBLAS_B200_API int32_t sblas_host16_num_cols(const sblas_host_matrix16 *h);
SBLAS_B200_API int64_t sblas_host16_entries(const sblas_host_matrix16 *h);
SBLAS_B200_API int32_t sblas_host16_nblocks(const sblas_host_matrix16 *h);
SBLAS_B200_API int32_t sblas_host16_table_size(const sblas_host_matrix16 *h);
SBLAS_B200_API void sblas_host16_members(const sblas_host_matrix16 *h, const uint16_t **pos_index, const uint8_t **val_index,
                                         const float **val_table, const int32_t **block_bounds,
                                         const int32_t **block_index_bounds);
SBLAS_B200_API sblas_b200_matrix *sblas_host16_device_matrix(sblas_host_matrix16 *h);

#ifdef __cplusplus
}
#endif
#endif
