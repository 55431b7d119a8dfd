"""The oracle pin, repeated where the GPU parity tests run ("what's weak" 2).

The GPU parity tests compare the CUDA path with the C port (oracle/liboracle.so).  The port <-> reference pin lives in
tests/test_oracle_pinned.py, which is `-m "not gpu"` and therefore not part of the run on the GPU box.  These tests carry the
`gpu` mark only so that they run THERE: they check that the unmodified reference library shipped to this box
(oracle/_ref/libsparsematrix_ref.so, built from /root/reference by oracle/build_ref.sh) is present, passes its own SelfTest
(reference src/sparse/sparse-matrix.cc:209-313), and agrees bit for bit with the port on a reduced case list -- and that the
CUDA path agrees with the reference library itself, not only with the port.  Collected late: a missing _ref fails here
without hiding the parity tests before it.
"""
import numpy as np
import pytest

import sparsematrix_b200 as sb
from sparsematrix_b200 import gen
from oracle import oracle as O
from helpers import assert_bits_equal

pytestmark = pytest.mark.gpu

CASES = [
    # n (W rows), k, density, T, m, alpha, beta
    (300, 257, 0.1, 255, 7, 1.3, 2.0),
    (1024, 512, 0.01, 255, 31, 1.0, 0.0),          # fillers
    (1023, 2047, 0.25, 255, 9, 1.0, 1.0),          # harness shape (blas_test.cc:32)
]


def test_reference_library_is_on_this_box_and_passes_its_selftest():
    assert O.have_ref(), "oracle/_ref/libsparsematrix_ref.so did not travel to this box"
    assert O.ref_selftest()


@pytest.mark.parametrize("case", CASES)
def test_port_equals_reference_and_gpu_matches_both(case):
    assert O.have_ref(), "oracle/_ref/libsparsematrix_ref.so did not travel to this box"
    n, k, dens, T, m, alpha, beta = case
    codes = gen.codes_bernoulli(n, k, dens, seed=n + k, T=T)
    tab = gen.codebook(T, 1)
    R = O.RefMatrix()
    R.CopyForm(codes, n, k, k, tab, T, True)
    s = O.copyform(codes, n, k, k, tab, T, True)
    assert s.same_as(R.stream())
    assert_bits_equal(O.copyto(s, n, False), R.CopyTo(n, False), "CopyTo")
    a = gen.dense_uniform((m, k), 2)
    c0 = gen.dense_uniform((m, n), 3)
    c_ref = R.AddMatMat(a.copy(), m, k, c0.copy(), n, alpha, beta)
    assert_bits_equal(O.addmatmat(s, a, m, k, c0.copy(), n, alpha, beta), c_ref, "AddMatMat port vs reference")
    # the CUDA path against the reference library's own result
    S = sb.SparseMatrix(codes, n, k, k, tab, T, sb.SparseMatrix.Trans)
    got = S.AddMatMat(a, m, k, c0.copy(), n, alpha, beta)
    y64, bound = O.addmatmat_f64(s, a, m, k, c0, n, alpha, beta)
    back, _, _ = O.error_measures(got, m, n, n, y64, bound)
    rback, _, _ = O.error_measures(c_ref, m, n, n, y64, bound)
    assert back <= 1e-5, f"GPU backward error {back:.3e} (reference library {rback:.3e})"
    diff = np.abs(got.astype(np.float64) - c_ref.astype(np.float64))
    assert float((diff / np.maximum(bound, 1e-300)).max()) <= 2e-5
    assert_bits_equal(S.device_matrix().decode_dense(n, False), R.CopyTo(n, False), "device decode vs reference CopyTo")
