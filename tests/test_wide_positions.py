"""SparseMatrix<uint16, uint8, float>: two-byte position deltas.

The reference carries the code path (sparse-matrix.cc:24: zero_pad_interval = 2^(8 sizeof(PosIndex_t)) - 1) but instantiates
only <uint8, uint8, float> (:315).  The oracle for it is the UNMODIFIED reference instantiated with uint16 positions
(oracle/ref_shim16.cc includes the reference's .cc files and adds the instantiations) -> oracle/_ref/libsparsematrix_ref16.so;
the C port (sbo_*16) is pinned to it bit for bit here, and the CUDA path is checked against both."""
import numpy as np
import pytest

import sparsematrix_b200 as sb
from sparsematrix_b200 import gen
from oracle import oracle as O
from helpers import assert_bits_equal

needs_ref16 = pytest.mark.skipif(not O.have_ref16(), reason="oracle/_ref/libsparsematrix_ref16.so not built (needs /root/reference)")

CASES = [
    # n (W rows), k, density, T, trans
    (300, 257, 0.1, 255, True),
    (512, 4096, 0.001, 255, True),          # 99.9 % sparse: the 8-bit stream is filler dominated (8.9 B per non-zero)
    (1024, 2048, 0.01, 255, True),          # 99 %
    (257, 700, 0.02, 17, False),
    (300, 100000, 0.00002, 255, True),      # gaps beyond 65 535: fillers in the wide stream too
]


@needs_ref16
@pytest.mark.parametrize("case", CASES)
def test_wide_port_equals_reference_instantiated_with_uint16(case):
    n, k, dens, T, trans = case
    codes = gen.codes_bernoulli(n, k, dens, seed=n + k, T=T)
    tab = gen.codebook(T, 1)
    if trans:
        rows, cols, stride = n, k, k
    else:
        codes = np.ascontiguousarray(codes.T)
        rows, cols, stride = k, n, n
    R = O.Ref16Matrix()
    R.CopyForm(codes, rows, cols, stride, tab, T, trans)
    s = O.copyform(codes, rows, cols, stride, tab, T, trans, wide=True)
    assert s.pos.dtype == np.uint16 and s.same_as(R.stream())
    narrow = O.copyform(codes, rows, cols, stride, tab, T, trans)
    assert s.entries <= narrow.entries
    nnz = int((codes < T).sum())
    if case[2] >= 0.001:
        assert s.entries == nnz                                        # no fillers at these gaps
    else:
        assert s.entries > nnz and int(s.pos.max()) == 65535           # gaps beyond 65 535 are bridged (sparse-matrix.cc:46-51)
    assert_bits_equal(O.copyto(s, n + 2, False), R.CopyTo(n + 2, False), "CopyTo NoTrans")
    assert_bits_equal(O.copyto(s, k + 1, True), R.CopyTo(k + 1, True), "CopyTo Trans")
    m = 5
    a = gen.dense_uniform((m, k), 2); c0 = gen.dense_uniform((m, n), 3)
    c_ref = R.AddMatMat(a.copy(), m, k, c0.copy(), n, 1.3, 2.0)
    row_ptr, col, code, fill = O.stream_to_csr(s)
    assert fill == s.entries - nnz
    c_csr = O.addmatmat_csr(row_ptr, col, code, s.table, s.T, n, k, a, m, k, c0.copy(), n, 1.3, 2.0)
    assert_bits_equal(c_csr, c_ref, "AddMatMat: CSR restatement of the wide stream vs the reference<uint16>")
    # and the decoded matrix is the one the 8-bit stream holds
    assert_bits_equal(O.copyto(narrow, n, False), O.copyto(s, n, False), "same matrix in both widths")


def test_wide_host_class_builds_the_oracles_stream():
    """CopyForm of the host class for <uint16, uint8, float> (host scan): members bit-identical to the port; no GPU needed."""
    n, k, T = 512, 4096, 255
    codes = gen.codes_bernoulli(n, k, 0.001, seed=5)
    tab = gen.codebook(T, 1)
    S = sb.SparseMatrix16(codes, n, k, k, tab, T, sb.SparseMatrix16.Trans)
    s = O.copyform(codes, n, k, k, tab, T, True, wide=True)
    pos, code, table, Tm, bb, bib = S.members()
    assert pos.dtype == np.uint16 and np.array_equal(pos, s.pos) and np.array_equal(code, s.code)
    assert Tm == T and np.array_equal(bb, s.block_bounds) and np.array_equal(bib, s.block_index_bounds)
    assert (S.NumRows(), S.NumCols()) == (k, n)
    assert_bits_equal(S.CopyTo(n, sb.SparseMatrix16.NoTrans), O.copyto(s, n, False), "CopyTo")
    assert_bits_equal(S.CopyTo(k, sb.SparseMatrix16.Trans), O.copyto(s, k, True), "CopyTo Trans")


@pytest.mark.gpu
@pytest.mark.parametrize("n,k,dens", [(512, 4096, 0.001), (1024, 2048, 0.01), (700, 520, 0.08), (300, 100000, 0.00002)])
def test_wide_positions_through_the_device_path(n, k, dens):
    T = 255
    codes = gen.codes_bernoulli(n, k, dens, seed=n + k)
    tab = gen.codebook(T, 1)
    S = sb.SparseMatrix16(codes, n, k, k, tab, T, sb.SparseMatrix16.Trans)
    s = O.copyform(codes, n, k, k, tab, T, True, wide=True)
    dev = S.device_matrix()
    inf = dev.info()
    assert inf.nnz == int((codes < T).sum()) and inf.entries == s.entries and inf.fillers == s.entries - inf.nnz
    # structure: the device CSR and the device decode, bit for bit against the oracle's
    row_ptr, col, code, _ = O.stream_to_csr(s)
    d_rp, d_col, d_code = dev.export_csr()
    assert np.array_equal(d_rp, row_ptr) and np.array_equal(d_col, col) and np.array_equal(d_code, code)
    assert_bits_equal(dev.decode_dense(n, False), O.copyto(s, n, False), "device decode")
    # the same device matrix as the 8-bit stream of the same codes gives
    S8 = sb.SparseMatrix(codes, n, k, k, tab, T, sb.SparseMatrix.Trans)
    for m, alpha, beta in ((1, 1.0, 1.0), (7, 1.3, 2.0), (40, 1.0, 0.0)):
        a = gen.dense_uniform((m, k), 10 + m); c0 = gen.dense_uniform((m, n), 20 + m)
        got = S.AddMatMat(a, m, k, c0.copy(), n, alpha, beta)
        y64, bound = O.addmatmat_csr_f64(row_ptr, col, code, s.table, s.T, n, k, a, m, k, c0, n, alpha, beta)
        back = O.error_measures(got, m, n, n, y64, bound)[0]
        assert back <= 1e-5, (m, back)
        assert_bits_equal(got, S8.AddMatMat(a, m, k, c0.copy(), n, alpha, beta), "wide vs narrow stream of the same matrix")
    if O.have_ref16():
        R = O.Ref16Matrix()
        R.CopyForm(codes, n, k, k, tab, T, True)
        a = gen.dense_uniform((3, k), 1); c0 = gen.dense_uniform((3, n), 2)
        want = R.AddMatMat(a.copy(), 3, k, c0.copy(), n, 1.3, 2.0)
        got = S.AddMatMat(a, 3, k, c0.copy(), n, 1.3, 2.0)
        y64, bound = O.addmatmat_csr_f64(row_ptr, col, code, s.table, s.T, n, k, a, 3, k, c0, n, 1.3, 2.0)
        assert float((np.abs(got.astype(np.float64) - want) / np.maximum(bound, 1e-300)).max()) <= 2e-5
