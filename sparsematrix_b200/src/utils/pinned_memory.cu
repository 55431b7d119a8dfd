// pinned_memory.cu -- host buffers for the caller's A and C that the host entry point can overlap.
//
// sblas_b200_addmatmat_host() takes any host pointer (the reference's AddMatMat signature,
// sparse-matrix.cc:139), but only page-locked, device-mapped memory lets it run the copies beside
// the product and write the result straight into the caller's C (src/sparse/product.cu,
// host_product_overlapped).  These two calls hand out exactly that kind of memory, so a caller
// written against the reference needs no CUDA header to get the fast path.
#include "device_matrix.h"

using namespace sblas_b200;

extern "C" {

int sblas_b200_host_alloc(int64_t bytes, void **ptr) {
    if (!ptr) return fail(SBLAS_B200_EINVAL, "host_alloc: null output pointer");
    if (bytes <= 0) return fail(SBLAS_B200_EINVAL, "host_alloc: bytes = %lld", (long long)bytes);
    void *p = nullptr;
    const cudaError_t e = cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocMapped | cudaHostAllocPortable);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        return fail(SBLAS_B200_ENOMEM, "host_alloc: %lld bytes of page-locked memory", (long long)bytes);
    }
    if (e != cudaSuccess) return cuda_fail(e, "cudaHostAlloc", __FILE__, __LINE__);   // no device, no driver: no fallback to malloc
    *ptr = p;
    return SBLAS_B200_OK;
}

int sblas_b200_host_free(void *ptr) {
    if (!ptr) return SBLAS_B200_OK;
    SB_CUDA(cudaFreeHost(ptr));
    return SBLAS_B200_OK;
}

}  // extern "C"
