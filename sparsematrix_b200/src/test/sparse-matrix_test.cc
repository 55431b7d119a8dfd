// sparse-matrix_test.cc -- the reference's unit test main (src/sparse/sparse-matrix_test.cc there):
// run SelfTest, which here goes through the device path.  Exit code != 0 on failure.
#include <iostream>

#include "sparse-matrix.h"

int main() {
    sblas::SparseMatrix<uint8, uint8, float> matrix;
    const bool ok = matrix.SelfTest();
    std::cout << (ok ? "SparseMatrix Test success..." : "SparseMatrix Test fail...") << std::endl;
    return ok ? 0 : -1;
}
