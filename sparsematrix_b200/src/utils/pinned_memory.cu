// pinned_memory.cu -- host buffers for the caller's A and C that the host entry point can overlap.
//
// sblas_b200_addmatmat_host() takes any host pointer (the reference's AddMatMat signature,
// sparse-matrix.cc:139), but only page-locked, device-mapped memory lets it run the copies beside
// the product and write the result straight into the caller's C (src/sparse/product.cu,
// host_product_overlapped).  These two calls hand out exactly that kind of memory, so a caller
// written against the reference needs no CUDA header to get the fast path.
#include <stdlib.h>

#include <map>

#include "device_matrix.h"

using namespace sblas_b200;

// ---- registration cache for pageable caller buffers -------------------------------------------
// An unmodified caller of the reference hands AddMatMat plain malloc'ed memory (blas_test.h:121).  Copies from
// pageable memory are staged by the driver at ~10 GB/s and cannot overlap the product; pinning a buffer per call
// costs >= 0.9 ms (profiles/r01d_e2e.md).  So: a buffer the host entry point sees a SECOND time at the same
// address with the same size is registered once (cudaHostRegister, mapped + portable) and stays registered; from
// then on the call takes the overlapped route.  Owner = the handle whose call registered it: sblas_b200_free()
// of that handle unregisters, sblas_b200_host_forget() does it on request, and at most kRegisterCapBytes stay
// registered (oldest first out).  Opt-in: SBLAS_B200_HOST_REGISTER=1 (see register_enabled()).
namespace sblas_b200 {

namespace {
struct HostRange {
    size_t bytes = 0;
    int seen = 0;
    bool registered = false, refused = false;
    const void *owner = nullptr;
    uint64_t stamp = 0;
};
std::mutex &reg_lock() { static std::mutex *m = new std::mutex; return *m; }
std::map<uintptr_t, HostRange> &reg_table() { static auto *t = new std::map<uintptr_t, HostRange>; return *t; }
size_t g_registered_bytes = 0;
uint64_t g_stamp = 0;
constexpr size_t kRegisterCapBytes = (size_t)4 << 30;
constexpr size_t kRegisterMinBytes = (size_t)64 << 10;     // below this the staged copy is as fast as anything

// Opt-in (SBLAS_B200_HOST_REGISTER=1, read per call).  Default off: a registration outlives free() of the block, and a
// caller that frees and reallocates large buffers between calls (numpy does, all the time) would have its new block
// read through the old pages -- tests/test_gpu_parity.py::test_powerlaw_spmv_with_long_rows tripped over exactly that
// when this was on by default.  Callers that keep their buffers switch it on.
bool register_enabled() {
    const char *e = getenv("SBLAS_B200_HOST_REGISTER");
    return e && e[0] == '1';
}

void drop_locked(std::map<uintptr_t, HostRange>::iterator it) {
    if (it->second.registered) {
        if (cudaHostUnregister(reinterpret_cast<void *>(it->first)) != cudaSuccess) cudaGetLastError();
        g_registered_bytes -= it->second.bytes;
    }
    reg_table().erase(it);
}
}  // namespace

// Called by the host entry point for each caller buffer before it asks whether the buffer is mapped.
void note_host_buffer(const void *owner, const void *ptr, size_t bytes) {
    if (!register_enabled() || !ptr || bytes < kRegisterMinBytes) return;
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, ptr) != cudaSuccess) { cudaGetLastError(); return; }
    if (attr.type != cudaMemoryTypeUnregistered) return;    // already page-locked (ours or the caller's), or not host memory
    std::lock_guard<std::mutex> guard(reg_lock());
    auto &t = reg_table();
    const uintptr_t key = reinterpret_cast<uintptr_t>(ptr);
    auto it = t.find(key);
    if (it != t.end() && it->second.bytes != bytes) { drop_locked(it); it = t.end(); }   // same address, other size: a new buffer
    if (it == t.end()) {
        if (t.size() >= 4096) return;                       // a caller that never reuses buffers: stop remembering
        HostRange r;
        r.bytes = bytes; r.seen = 1; r.owner = owner; r.stamp = ++g_stamp;
        t[key] = r;
        return;
    }
    HostRange &r = it->second;
    r.stamp = ++g_stamp;
    if (r.refused) return;
    if (++r.seen < 2) return;
    while (g_registered_bytes + bytes > kRegisterCapBytes) {        // make room: oldest registered range out
        auto victim = t.end();
        for (auto jt = t.begin(); jt != t.end(); ++jt)
            if (jt->second.registered && jt != it && (victim == t.end() || jt->second.stamp < victim->second.stamp)) victim = jt;
        if (victim == t.end()) break;
        drop_locked(victim);
    }
    if (g_registered_bytes + bytes > kRegisterCapBytes) { r.refused = true; return; }
    const cudaError_t e = cudaHostRegister(const_cast<void *>(ptr), bytes, cudaHostRegisterMapped | cudaHostRegisterPortable);
    if (e != cudaSuccess) { cudaGetLastError(); r.refused = true; return; }     // overlaps a foreign registration, read-only pages, ...
    r.registered = true;
    r.owner = owner;
    g_registered_bytes += bytes;
}

void forget_host_buffers(const void *owner, const void *ptr) {
    std::lock_guard<std::mutex> guard(reg_lock());
    auto &t = reg_table();
    for (auto it = t.begin(); it != t.end();) {
        auto next = std::next(it);
        const bool hit = ptr ? it->first == reinterpret_cast<uintptr_t>(ptr) : (!owner || it->second.owner == owner);
        if (hit) drop_locked(it);
        it = next;
    }
}

}  // namespace sblas_b200

extern "C" {

int sblas_b200_host_forget(const void *ptr) {
    forget_host_buffers(nullptr, ptr);
    return SBLAS_B200_OK;
}

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
