// sparse-matrix.h -- host-side class of the B200 build.
//
// Mirrors the public interface of the reference class
// sblas::SparseMatrix<PosIndex_t, ValIndex_t, Value_t, block_row_shift, block_col_shift>
// (reference src/sparse/sparse-matrix.h:25-53): same names, same argument meaning,
// same data members in the same order (so objects are layout compatible with code
// built against the reference header), same "print and return" error convention.
// What differs is behind AddMatMat: instead of driving the CPU micro-kernels of
// kernel.cc it forwards to the CUDA library through the C ABI in
// include/sblas_b200.h.  The compressed members are uploaded once per object
// (a registry keyed by the object's address, src/utils/device_registry.h) and the
// device copy is dropped by Destroy(), which CopyForm and the destructor call --
// exactly where the reference clears its vectors (sparse-matrix.cc:10-18,22).
//
// There is no CPU fallback: AddMatMat needs a CUDA device.
#ifndef SBLAS_B200_SPARSE_MATRIX_H_
#define SBLAS_B200_SPARSE_MATRIX_H_

#include <stdint.h>

#include <utility>
#include <vector>

typedef int8_t int8;
typedef int16_t int16;
typedef int32_t int32;
typedef uint8_t uint8;
typedef uint16_t uint16;
typedef uint32_t uint32;

// Panel geometry, the reference's compile-time constants (kernel.h:25-26).
#ifndef SBLAS_BLOCK_ROW_SHIFT
#define SBLAS_BLOCK_ROW_SHIFT 0
#endif
#ifndef SBLAS_BLOCK_COL_SHIFT
#define SBLAS_BLOCK_COL_SHIFT 8
#endif

struct sblas_b200_matrix;

namespace sblas {

enum SBLAS_TRANSPOSE { SblasNoTrans = 0, SblasTrans = 1 };

template <typename PosIndex_t, typename ValIndex_t, typename Value_t,
          const int32 block_row_shift = SBLAS_BLOCK_ROW_SHIFT, const int32 block_col_shift = SBLAS_BLOCK_COL_SHIFT>
class SparseMatrix {
public:
    SparseMatrix() : rows_(0), cols_(0) {}
    // density_matrix holds codebook indices; an element is stored iff its index < val_table_size.
    SparseMatrix(const ValIndex_t *density_matrix, int32 rows, int32 cols, int32 stride, const Value_t *vals,
                 int32 val_table_size, SBLAS_TRANSPOSE trans = SblasNoTrans)
        : rows_(0), cols_(0) {
        CopyForm(density_matrix, rows, cols, stride, vals, val_table_size, trans);
    }
    ~SparseMatrix() { Destroy(); }
    // Copies take the compressed members only.  The device copy is registered per object address
    // (src/utils/device_registry.h): an assigned-to object drops its own -- it would describe the old
    // contents -- and uploads again on its next AddMatMat.  Member functions: no effect on layout.
    SparseMatrix(const SparseMatrix &oth)
        : pos_index_(oth.pos_index_), val_index_(oth.val_index_), val_table_(oth.val_table_), block_bounds_(oth.block_bounds_),
          block_index_bounds_(oth.block_index_bounds_), rows_(oth.rows_), cols_(oth.cols_) {}
    SparseMatrix &operator=(const SparseMatrix &oth);

    // Clears the compressed members and drops the device copy.
    void Destroy();

    // Build the compressed form from a dense index matrix (rows x cols, leading dimension stride).
    // SblasTrans stores the transpose: NumRows() == cols, NumCols() == rows afterwards.
    void CopyForm(const ValIndex_t *density_matrix, int32 rows, int32 cols, int32 stride, const Value_t *vals,
                  int32 val_table_size, SBLAS_TRANSPOSE trans = SblasNoTrans);

    // Expand to dense values (NumRows() x stride, or NumCols() x stride when trans).
    void CopyTo(Value_t *density_matrix, int32 stride, SBLAS_TRANSPOSE trans = SblasNoTrans);

    // C[m x NumCols()] = beta*C + alpha * A[m x NumRows()] * (*this); host pointers, row-major.
    void AddMatMat(Value_t *a, int32 m, int32 lda, Value_t *c, int32 ldc, Value_t alpha, Value_t beta);

    int32 NumRows() const { return rows_; }
    int32 NumCols() const { return cols_; }

    bool operator==(const SparseMatrix<PosIndex_t, ValIndex_t, Value_t, block_row_shift, block_col_shift> &oth);
    bool SelfTest();

    // B200 additions (no effect on layout): the uploaded device matrix of this object on the
    // current CUDA device (uploading on first use; nullptr + message on failure), for callers
    // that keep A and C on the device and call sblas_b200_addmatmat_dev themselves.
    sblas_b200_matrix *DeviceMatrix();

private:
    std::vector<PosIndex_t> pos_index_;                        // delta of (row*256 + col) inside a panel
    std::vector<ValIndex_t> val_index_;                        // codebook index; == table size marks a filler
    std::vector<Value_t> val_table_;                           // codebook + one trailing zero
    std::vector<std::pair<int32, int32> > block_bounds_;       // (row_off, col_off) of each non-empty panel
    std::vector<std::pair<int32, int32> > block_index_bounds_; // [start, end) of the panel in the streams
    int32 rows_;
    int32 cols_;
};

}  // namespace sblas

#endif
