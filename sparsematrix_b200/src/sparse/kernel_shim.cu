// kernel_shim.cu -- the reference's FREE kernel functions (src/sparse/kernel.h:42-62 there) on the device,
// for callers that use them directly instead of going through SparseMatrix::AddMatMat:
//   sblas_beta_operation_kernel     kernel.cc:10-29    -> sblas_b200_scale_host
//   sblas_trans_kernel              kernel.cc:31-187   -> sblas_b200_transpose_host
//   sblas_kernel_operation{,_naive} kernel.cc:213-369  -> sblas_b200_panel_product_host(trans = 0)
//   sblas_kernel_operation_trans{,_ex} kernel.cc:612-800 -> sblas_b200_panel_product_host(trans = 1)
// Host pointers in and out, synchronous, like the functions they stand in for.  The panel product
// uploads the one panel it is given (stream -> CSR on the device) and runs the ordinary product
// kernels on it, so it is the same arithmetic as AddMatMat; it is a convenience, not a fast path
// (AddMatMat keeps the matrix resident).
#include <algorithm>

#include "device_matrix.h"

namespace sblas_b200 {
namespace {

// out[j * ld_out + i] = in[i * ld_in + j], i < rows, j < cols
__global__ void __launch_bounds__(256)
transpose_kernel(const float *__restrict__ in, int rows, int cols, int ld_in, float *__restrict__ out, int ld_out) {
    __shared__ float tile[32][33];
    const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int ii = ty; ii < 32; ii += 8) {
        const int i = i0 + ii, j = j0 + tx;
        if (i < rows && j < cols) tile[ii][tx] = in[(size_t)i * ld_in + j];
    }
    __syncthreads();
    for (int jj = ty; jj < 32; jj += 8) {
        const int j = j0 + jj, i = i0 + tx;
        if (i < rows && j < cols) out[(size_t)j * ld_out + i] = tile[tx][jj];
    }
}

__global__ void __launch_bounds__(256)
scale_rows_kernel(float *__restrict__ c, int m, int n, float beta) {
    const size_t total = (size_t)m * n;
    for (size_t i = blockIdx.x * (size_t)256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) c[i] *= beta;
}

struct DeviceBuffers {                                   // frees whatever was allocated when the call ends
    float *p[3] = {nullptr, nullptr, nullptr};
    ~DeviceBuffers() { for (float *q : p) cudaFree(q); }
};

int current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return -1; }
    return dev;
}

}  // namespace
}  // namespace sblas_b200

using namespace sblas_b200;

extern "C" {

int sblas_b200_scale_host(float *c, int32_t m, int32_t n, int32_t ldc, float beta) {
    if (m < 0 || n < 0) return fail(SBLAS_B200_EINVAL, "scale: negative size");
    if (m == 0 || n == 0 || beta == 1.0f) return SBLAS_B200_OK;          // kernel.cc:10-29 is only called when beta != 1
    if (!c || ldc < n) return fail(SBLAS_B200_EINVAL, "scale: c is null or ldc %d < n %d", ldc, n);
    DeviceBuffers b;
    SB_CUDA(cudaMalloc((void **)&b.p[0], (size_t)m * n * 4));
    SB_CUDA(cudaMemcpy2D(b.p[0], (size_t)n * 4, c, (size_t)ldc * 4, (size_t)n * 4, m, cudaMemcpyHostToDevice));
    SB_LAUNCH(scale_rows_kernel, (unsigned)std::min<size_t>(((size_t)m * n + 255) / 256, 148 * 8), 256, 0, 0, b.p[0], m, n, beta);
    SB_CHECK_LAUNCH();
    SB_CUDA(cudaMemcpy2D(c, (size_t)ldc * 4, b.p[0], (size_t)n * 4, (size_t)n * 4, m, cudaMemcpyDeviceToHost));
    return SBLAS_B200_OK;
}

int sblas_b200_transpose_host(const float *a, int32_t m, int32_t n, int32_t lda, float *sa, int32_t ldsa) {
    if (m < 0 || n < 0) return fail(SBLAS_B200_EINVAL, "transpose: negative size");
    if (m == 0 || n == 0) return SBLAS_B200_OK;
    if (!a || !sa || lda < n || ldsa < m) return fail(SBLAS_B200_EINVAL, "transpose: null pointer or leading dimension too small");
    DeviceBuffers b;
    SB_CUDA(cudaMalloc((void **)&b.p[0], (size_t)m * n * 4));
    SB_CUDA(cudaMalloc((void **)&b.p[1], (size_t)m * n * 4));
    SB_CUDA(cudaMemcpy2D(b.p[0], (size_t)n * 4, a, (size_t)lda * 4, (size_t)n * 4, m, cudaMemcpyHostToDevice));
    SB_LAUNCH(transpose_kernel, dim3((n + 31) / 32, (m + 31) / 32), 256, 0, 0, b.p[0], m, n, n, b.p[1], m);
    SB_CHECK_LAUNCH();
    // only the n x m result is written: what lies beyond m in a row of sa stays the caller's
    SB_CUDA(cudaMemcpy2D(sa, (size_t)ldsa * 4, b.p[1], (size_t)m * 4, (size_t)m * 4, n, cudaMemcpyDeviceToHost));
    return SBLAS_B200_OK;
}

int sblas_b200_panel_product_host(int32_t trans, int32_t m, int32_t n, int32_t k, const float *a, int32_t lda, float *c,
                                  int32_t ldc, float alpha, const uint8_t *ppos, const uint8_t *pval, int32_t pos_len,
                                  const float *val_table, int32_t valid_table_size) {
    if (m < 0 || n < 0 || k < 0 || pos_len < 0) return fail(SBLAS_B200_EINVAL, "panel product: negative size");
    if (n > kPanelWidth) return fail(SBLAS_B200_EINVAL, "panel product: n = %d exceeds the panel width %d", n, kPanelWidth);
    if (m == 0 || n == 0 || pos_len == 0 || alpha == 0.0f) return SBLAS_B200_OK;
    if (!a || !c) return fail(SBLAS_B200_EINVAL, "panel product: null operand");
    const int dev = current_device();
    if (dev < 0) return fail(SBLAS_B200_ECUDA, "panel product: no CUDA device");
    const int32_t bb[2] = {0, 0}, bib[2] = {0, pos_len};
    sblas_b200_matrix *mat = nullptr;
    int rc = sblas_b200_upload(ppos, pval, pos_len, val_table, valid_table_size, bb, bib, 1, k, n, dev, &mat);
    if (rc != SBLAS_B200_OK) return rc;
    if (!trans) {                                        // a: m x k (lda), c: m x n (ldc), row-major
        rc = sblas_b200_addmatmat_host(mat, a, m, lda, c, ldc, alpha, 1.0f);
        sblas_b200_free(mat);
        return rc;
    }
    // transposed operands (sparse-matrix.cc:181-189): a is k x lda with the m values of one S row contiguous,
    // c is n x ldc likewise.  Bring both in, turn them on the device, multiply, turn C back.
    if (lda < m || ldc < m) { sblas_b200_free(mat); return fail(SBLAS_B200_EINVAL, "panel product: leading dimension below m"); }
    DeviceBuffers b;
    auto done = [&](int code) { sblas_b200_free(mat); return code; };
    cudaError_t e = cudaMalloc((void **)&b.p[0], (size_t)std::max(k, n) * m * 4);        // staging of a^T / c^T as given
    if (e == cudaSuccess) e = cudaMalloc((void **)&b.p[1], (size_t)m * k * 4);           // A  m x k
    if (e == cudaSuccess) e = cudaMalloc((void **)&b.p[2], (size_t)m * n * 4);           // C  m x n
    if (e == cudaSuccess) e = cudaMemcpy2D(b.p[0], (size_t)m * 4, a, (size_t)lda * 4, (size_t)m * 4, k, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return done(cuda_fail(e, "panel product: staging", __FILE__, __LINE__));
    SB_LAUNCH(transpose_kernel, dim3((m + 31) / 32, (k + 31) / 32), 256, 0, 0, b.p[0], k, m, m, b.p[1], k);
    e = cudaMemcpy2D(b.p[0], (size_t)m * 4, c, (size_t)ldc * 4, (size_t)m * 4, n, cudaMemcpyHostToDevice);   // (default stream: after the kernel)
    if (e != cudaSuccess) return done(cuda_fail(e, "panel product: staging c", __FILE__, __LINE__));
    SB_LAUNCH(transpose_kernel, dim3((m + 31) / 32, (n + 31) / 32), 256, 0, 0, b.p[0], n, m, m, b.p[2], n);
    rc = sblas_b200_addmatmat_dev(mat, b.p[1], m, k, b.p[2], n, alpha, 1.0f, nullptr);
    if (rc != SBLAS_B200_OK) { cudaDeviceSynchronize(); return done(rc); }
    SB_LAUNCH(transpose_kernel, dim3((n + 31) / 32, (m + 31) / 32), 256, 0, 0, b.p[2], m, n, n, b.p[0], m);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy2D(c, (size_t)ldc * 4, b.p[0], (size_t)m * 4, (size_t)m * 4, n, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return done(cuda_fail(e, "panel product: result", __FILE__, __LINE__));
    return done(SBLAS_B200_OK);
}

}  // extern "C"
