// device_registry.h -- side table "host object -> uploaded device matrix".
//
// The reference class has private members and no spare field (sparse-matrix.h:45-52),
// so the device copy cannot live inside the object without changing its layout.  It
// lives here instead, keyed by the object's address; Destroy() (called by CopyForm and
// the destructor) evicts.  A copied object simply misses and uploads again.
#ifndef SBLAS_B200_DEVICE_REGISTRY_H_
#define SBLAS_B200_DEVICE_REGISTRY_H_

#include <stdint.h>

struct sblas_b200_matrix;

namespace sblas {
namespace device {

struct StreamView {                 // the compressed members, as plain pointers
    const void *pos_index;          // pos_bytes per entry
    int32_t pos_bytes;              // sizeof(PosIndex_t): 1 or 2
    const uint8_t *val_index;
    int64_t entries;
    const float *val_table;
    int32_t valid_table_size;
    const int32_t *block_bounds;        // pairs
    const int32_t *block_index_bounds;  // pairs
    int32_t nblocks;
    int32_t rows, cols;
};

// Device matrix for `owner` on the current CUDA device, uploading `view` on a miss.
// Returns nullptr after printing the library's error message.
sblas_b200_matrix *Acquire(const void *owner, const StreamView &view);

// Register a device matrix built elsewhere (device CopyForm) for `owner` on device `dev`; the
// registry owns it from here on.  An existing entry is freed first.
void Adopt(const void *owner, int dev, sblas_b200_matrix *mat);

// Ordinal of the current CUDA device, -1 when there is none.
int CurrentDevice();

// Drop and free whatever is registered for `owner` (no-op when nothing is).
void Evict(const void *owner);

}  // namespace device
}  // namespace sblas

#endif
