// peer_memory.cu -- device allocations another process (one process per GPU) can map, so that a
// rank's product kernels write their slab of C straight into the memory of the GPU that wants the
// whole result ("kernel epilogue stores straight into peer-mapped Y").
//
// The reference has no counterpart: its panels are independent (sparse-matrix.cc:164-190) and all
// live in one address space.  Here the output rows of a slab are independent in the same way, so
// "reassembly" needs no collective at all -- only a C pointer that happens to be NVLink-mapped.
// The product entry points take it like any other device pointer (c + col_begin, ldc = n).
//
// cudaIpc handles are the plumbing: 64 opaque bytes the owner hands to its peers through whatever
// channel the application has (torch.distributed in sparsematrix_b200/sharding.py).
#include <string.h>

#include <map>

#include "device_matrix.h"

namespace sblas_b200 {
namespace {
std::mutex g_peer_lock;
std::map<void *, int> g_peer_owned;    // pointer -> device, allocations of sblas_b200_peer_alloc
struct Mapping { size_t bytes; int device; };   // device < 0: a range of sblas_b200_peer_register (not ours to unmap)
std::map<uintptr_t, Mapping> g_peer_opened;   // base -> extent, mappings of sblas_b200_peer_open / _peer_register
std::atomic<int> g_peer_open_count{0};
}  // namespace

bool is_peer_pointer(const void *p) {
    if (g_peer_open_count.load(std::memory_order_relaxed) == 0) return false;   // the common case costs one load
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    std::lock_guard<std::mutex> guard(g_peer_lock);
    auto it = g_peer_opened.upper_bound(a);
    if (it == g_peer_opened.begin()) return false;
    --it;
    return a - it->first < it->second.bytes;
}
}  // namespace sblas_b200

using namespace sblas_b200;

extern "C" {

int sblas_b200_peer_alloc(int32_t device, int64_t bytes, void **d_ptr, uint8_t *handle) {
    if (!d_ptr || !handle) return fail(SBLAS_B200_EINVAL, "peer_alloc: null output pointer");
    if (bytes <= 0) return fail(SBLAS_B200_EINVAL, "peer_alloc: bytes = %lld", (long long)bytes);
    static_assert(sizeof(cudaIpcMemHandle_t) == SBLAS_B200_IPC_HANDLE_BYTES, "handle size is part of the ABI");
    DeviceScope scope(device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, (size_t)bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(SBLAS_B200_ENOMEM, "peer_alloc: cudaMalloc of %lld bytes: %s", (long long)bytes, cudaGetErrorString(e));
    }
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return cuda_fail(e, "cudaIpcGetMemHandle", __FILE__, __LINE__);
    }
    memcpy(handle, &h, sizeof(h));
    {
        std::lock_guard<std::mutex> guard(g_peer_lock);
        g_peer_owned[p] = device;
    }
    *d_ptr = p;
    return SBLAS_B200_OK;
}

int sblas_b200_peer_open(int32_t device, const uint8_t *handle, int64_t bytes, void **d_ptr) {
    if (!d_ptr || !handle) return fail(SBLAS_B200_EINVAL, "peer_open: null pointer");
    if (bytes <= 0) return fail(SBLAS_B200_EINVAL, "peer_open: bytes = %lld", (long long)bytes);
    DeviceScope scope(device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void *p = nullptr;
    SB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    {
        std::lock_guard<std::mutex> guard(g_peer_lock);
        g_peer_opened[reinterpret_cast<uintptr_t>(p)] = Mapping{(size_t)bytes, device};
        g_peer_open_count.fetch_add(1, std::memory_order_relaxed);
    }
    *d_ptr = p;
    return SBLAS_B200_OK;
}

int sblas_b200_peer_close(void *d_ptr) {
    int device;
    {
        std::lock_guard<std::mutex> guard(g_peer_lock);
        auto it = g_peer_opened.find(reinterpret_cast<uintptr_t>(d_ptr));
        if (it == g_peer_opened.end() || it->second.device < 0)
            return fail(SBLAS_B200_EINVAL, "peer_close: not a pointer of sblas_b200_peer_open");
        device = it->second.device;
        g_peer_opened.erase(it);
        g_peer_open_count.fetch_sub(1, std::memory_order_relaxed);
    }
    DeviceScope scope(device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    SB_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return SBLAS_B200_OK;
}

// Ranges of peer memory the application mapped itself (one process driving several GPUs with
// cudaDeviceEnablePeerAccess, cuMemMap of a fabric handle, a symmetric heap): the same epilogues apply.
int sblas_b200_peer_register(void *d_ptr, int64_t bytes) {
    if (!d_ptr) return fail(SBLAS_B200_EINVAL, "peer_register: null pointer");
    if (bytes <= 0) return fail(SBLAS_B200_EINVAL, "peer_register: bytes = %lld", (long long)bytes);
    const uintptr_t a = reinterpret_cast<uintptr_t>(d_ptr);
    std::lock_guard<std::mutex> guard(g_peer_lock);
    auto next = g_peer_opened.lower_bound(a);                   // first range starting at or after a
    if (next != g_peer_opened.end() && next->first - a < (uintptr_t)bytes)
        return fail(SBLAS_B200_EINVAL, "peer_register: overlaps a range already known");
    if (next != g_peer_opened.begin()) {
        auto prev = next;
        --prev;
        if (a - prev->first < prev->second.bytes) return fail(SBLAS_B200_EINVAL, "peer_register: overlaps a range already known");
    }
    g_peer_opened[a] = Mapping{(size_t)bytes, -1};
    g_peer_open_count.fetch_add(1, std::memory_order_relaxed);
    return SBLAS_B200_OK;
}

int sblas_b200_peer_unregister(void *d_ptr) {
    std::lock_guard<std::mutex> guard(g_peer_lock);
    auto it = g_peer_opened.find(reinterpret_cast<uintptr_t>(d_ptr));
    if (it == g_peer_opened.end() || it->second.device >= 0)
        return fail(SBLAS_B200_EINVAL, "peer_unregister: not a pointer of sblas_b200_peer_register");
    g_peer_opened.erase(it);
    g_peer_open_count.fetch_sub(1, std::memory_order_relaxed);
    return SBLAS_B200_OK;
}

int sblas_b200_peer_free(void *d_ptr) {
    int device;
    {
        std::lock_guard<std::mutex> guard(g_peer_lock);
        auto it = g_peer_owned.find(d_ptr);
        if (it == g_peer_owned.end()) return fail(SBLAS_B200_EINVAL, "peer_free: not a pointer of sblas_b200_peer_alloc");
        device = it->second;
        g_peer_owned.erase(it);
    }
    DeviceScope scope(device);
    if (!scope.ok()) return SBLAS_B200_ECUDA;
    SB_CUDA(cudaFree(d_ptr));
    return SBLAS_B200_OK;
}

}  // extern "C"
