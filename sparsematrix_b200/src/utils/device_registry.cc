#include "device_registry.h"

#include <cuda_runtime_api.h>
#include <stdio.h>

#include <map>
#include <mutex>
#include <utility>

#include "sblas_b200.h"

namespace sblas {
namespace device {

namespace {
// Both are heap objects that are never destroyed: a SparseMatrix with static storage duration may run its
// destructor (-> Evict) after the function-local statics of this file would have been torn down.
std::mutex &g_lock = *new std::mutex;
// (owner, device ordinal) -> handle
std::map<std::pair<const void *, int>, sblas_b200_matrix *> &table() {
    static std::map<std::pair<const void *, int>, sblas_b200_matrix *> *t =
        new std::map<std::pair<const void *, int>, sblas_b200_matrix *>;
    return *t;
}
}  // namespace

sblas_b200_matrix *Acquire(const void *owner, const StreamView &v) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        printf("sblas_b200: no CUDA device available (this build has no CPU fallback)\n");
        return nullptr;
    }
    std::lock_guard<std::mutex> guard(g_lock);
    auto key = std::make_pair(owner, dev);
    auto it = table().find(key);
    if (it != table().end()) return it->second;
    sblas_b200_matrix *mat = nullptr;
    const int rc = v.pos_bytes == 2
        ? sblas_b200_upload_slab16(static_cast<const uint16_t *>(v.pos_index), v.val_index, v.entries, v.val_table, v.valid_table_size,
                                   v.block_bounds, v.block_index_bounds, v.nblocks, v.rows, v.cols, 0, v.cols, dev, &mat)
        : sblas_b200_upload(static_cast<const uint8_t *>(v.pos_index), v.val_index, v.entries, v.val_table, v.valid_table_size,
                            v.block_bounds, v.block_index_bounds, v.nblocks, v.rows, v.cols, dev, &mat);
    if (rc != SBLAS_B200_OK) {
        printf("sblas_b200: device upload failed (%d): %s\n", rc, sblas_b200_last_error());
        return nullptr;
    }
    table()[key] = mat;
    return mat;
}

void Adopt(const void *owner, int dev, sblas_b200_matrix *mat) {
    std::lock_guard<std::mutex> guard(g_lock);
    auto key = std::make_pair(owner, dev);
    auto it = table().find(key);
    if (it != table().end()) sblas_b200_free(it->second);
    table()[key] = mat;
}

int CurrentDevice() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return -1; }
    return dev;
}

void Evict(const void *owner) {
    std::lock_guard<std::mutex> guard(g_lock);
    auto &t = table();
    for (auto it = t.lower_bound(std::make_pair(owner, -1)); it != t.end() && it->first.first == owner;) {
        sblas_b200_free(it->second);
        it = t.erase(it);
    }
}

}  // namespace device
}  // namespace sblas
