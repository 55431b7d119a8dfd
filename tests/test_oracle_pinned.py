"""Pin the oracle (oracle/sblas_oracle.c) before anything trusts it:
  1. against the reference's own known answers (SelfTest, sparse-matrix.cc:209-313),
  2. against the committed golden dumps produced by the unmodified reference,
  3. bit-for-bit against the live reference build (oracle/_ref) on seeded inputs.
CPU only."""
import numpy as np
import pytest

from oracle import oracle as O
from sparsematrix_b200 import gen
from helpers import assert_bits_equal, golden_names, golden_stream, load_golden

needs_ref = pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built (needs /root/reference)")

TAB8 = np.array([1.1, 2.2, 3.3, 4.4, 5.5, 6.6, 7.7, 8.8], np.float32)


# ---------------------------------------------------------------- 1. the reference's own known answers
def test_kat1_selftest_values():
    # sparse-matrix.cc:211-227
    s = O.copyform(np.array([0, 255, 255, 3, 7, 255], np.uint8), 3, 2, 2, TAB8, 8, False)
    assert (s.rows, s.cols) == (3, 2)
    # SURVEY appendix A worked example (dump of the reference): one filler, one real delta-255
    assert s.pos.tolist() == [0, 255, 2, 255] and s.code.tolist() == [0, 8, 3, 7]
    assert s.block_bounds.tolist() == [[0, 0]] and s.block_index_bounds.tolist() == [[0, 4]]
    assert s.table.shape == (9,) and s.table[8] == 0
    dense = O.copyto(s, 2, False)
    assert dense.reshape(-1).tolist() == np.array([1.1, 0, 0, 4.4, 8.8, 0], np.float32).tolist()
    c = np.array([4, 8], np.float32)
    O.addmatmat(s, np.array([3.1, 5, 7], np.float32), 1, 3, c, 2, 1.3, 2.0)
    assert abs(c[0] - 92.513) <= 1e-3 and abs(c[1] - 44.6) <= 1e-3          # :225-226


def test_kat2_selftest_values_transposed_input():
    # sparse-matrix.cc:228-246
    s = O.copyform(np.array([0, 255, 7, 255, 3, 255], np.uint8), 2, 3, 3, TAB8, 8, True)
    assert (s.rows, s.cols) == (3, 2)
    assert O.copyto(s, 2, False).reshape(-1).tolist() == np.array([1.1, 0, 0, 4.4, 8.8, 0], np.float32).tolist()
    assert O.copyto(s, 3, True).reshape(-1).tolist() == np.array([1.1, 0, 8.8, 0, 4.4, 0], np.float32).tolist()
    c = np.array([4, 8], np.float32)
    O.addmatmat(s, np.array([3.1, 5, 7], np.float32), 1, 3, c, 2, 1.3, 2.0)
    assert abs(c[0] - 92.513) <= 1e-3 and abs(c[1] - 44.6) <= 1e-3


@pytest.mark.parametrize("trans", [False, True])
def test_roundtrip_shape_of_selftest(trans):
    # sparse-matrix.cc:247-311: 1023 x 511 (stride 512), 25 % dense, 63-entry table, exact round trip
    m, n, stride, T = 1023, 511, 512, 63
    codes = np.full((m, stride), 255, np.uint8)
    codes[:, :n] = gen.codes_bernoulli(m, n, 0.25, seed=7, T=T)
    tab = gen.codebook(T, 3)
    want = np.where(codes[:, :n] < T, tab[np.minimum(codes[:, :n], T - 1)], np.float32(0))
    s = O.copyform(codes, m, n, stride, tab, T, trans)
    got = O.copyto(s, stride, trans)            # same orientation back, like :281 / :296
    assert_bits_equal(got[:, :n], want, "round trip")


def test_empty_table_gives_empty_object():
    # sparse-matrix.cc:26
    s = O.copyform(np.zeros(12, np.uint8), 3, 4, 4, np.zeros(1, np.float32), 0, False)
    assert (s.rows, s.cols, s.entries, s.nblocks) == (0, 0, 0, 0)


# ---------------------------------------------------------------- 2. golden dumps of the reference
@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_golden(name):
    g = load_golden(name)
    s = O.copyform(g["codes"], int(g["rows"]), int(g["cols"]), int(g["stride"]), g["vals"], int(g["T"]), bool(g["trans"]))
    assert s.same_as(golden_stream(g)), "compressed members differ from the reference dump"
    assert_bits_equal(O.copyto(s, int(g["stride_notrans"]), False), g["dense_notrans"], "CopyTo NoTrans")
    assert_bits_equal(O.copyto(s, int(g["stride_trans"]), True), g["dense_trans"], "CopyTo Trans")
    c = g["c_in"].copy()
    O.addmatmat(s, g["a"], int(g["m"]), int(g["lda"]), c, int(g["ldc"]), float(g["alpha"]), float(g["beta"]))
    assert_bits_equal(c, g["c_out"], "AddMatMat")
    # CSR restatement gives the same bits
    row_ptr, col, code, fill = O.stream_to_csr(s)
    assert fill == int((g["code"] >= int(g["T"])).sum())
    c2 = g["c_in"].copy()
    O.addmatmat_csr(row_ptr, col, code, s.table, s.T, s.cols, s.rows, g["a"], int(g["m"]), int(g["lda"]),
                    c2, int(g["ldc"]), float(g["alpha"]), float(g["beta"]))
    assert_bits_equal(c2, g["c_out"], "AddMatMat (CSR restatement)")


def test_golden_kat_matches_selftest_numbers():
    for name in ("kat1", "kat2"):
        g = load_golden(name)
        assert abs(g["c_out"][0] - 92.513) <= 1e-3 and abs(g["c_out"][1] - 44.6) <= 1e-3
        assert g["dense_notrans"].reshape(-1).tolist() == np.array([1.1, 0, 0, 4.4, 8.8, 0], np.float32).tolist()


# ---------------------------------------------------------------- 3. the live reference
@needs_ref
def test_reference_selftest_passes():
    assert O.ref_selftest()


CASES = [
    # n (W rows) , k, density, T, trans, m, lda_pad, ldc_pad, alpha, beta
    (64, 64, 0.5, 255, True, 1, 0, 0, 1.0, 1.0),
    (300, 257, 0.1, 255, True, 7, 3, 5, 1.3, 2.0),
    (257, 300, 0.1, 17, False, 8, 0, 0, 1.8, 1.2),
    (1023, 2047, 0.25, 255, True, 9, 0, 0, 1.0, 1.0),        # harness shape (blas_test.cc:32), small m
    (1024, 512, 0.01, 255, True, 31, 1, 1, 1.0, 0.0),         # fillers (7.7 % of gaps at 1 %)
    (512, 4096, 0.001, 255, True, 2, 0, 0, 1.3, 2.0),         # filler-dominated
    (256, 100, 1.0, 255, True, 3, 0, 0, 1.0, 1.0),            # fully dense, first delta 0
    (513, 64, 0.3, 1, False, 4, 0, 0, 0.0, 0.5),              # T=1, alpha=0
]


@needs_ref
@pytest.mark.parametrize("case", CASES)
def test_oracle_equals_reference(case):
    n, k, dens, T, trans, m, pa, pc, alpha, beta = case
    codes = gen.codes_bernoulli(n, k, dens, seed=n + k, T=T)
    tab = gen.codebook(T, 1)
    if trans:       # harness orientation: codes is n x k, S = codes^T (k x n)
        rows, cols, stride = n, k, k
    else:           # codes is S itself (k x n)
        codes = np.ascontiguousarray(codes.T)
        rows, cols, stride = k, n, n
    R = O.RefMatrix()
    R.CopyForm(codes, rows, cols, stride, tab, T, trans)
    s = O.copyform(codes, rows, cols, stride, tab, T, trans)
    assert (s.rows, s.cols) == (k, n)
    assert s.same_as(R.stream())
    assert_bits_equal(O.copyto(s, n + 2, False), R.CopyTo(n + 2, False), "CopyTo NoTrans")
    assert_bits_equal(O.copyto(s, k + 1, True), R.CopyTo(k + 1, True), "CopyTo Trans")
    lda, ldc = k + pa, n + pc
    a = gen.dense_uniform((m, lda), 2)
    c0 = gen.dense_uniform((m, ldc), 3)
    c_ref = R.AddMatMat(a.copy(), m, lda, c0.copy(), ldc, alpha, beta)
    c_orc = O.addmatmat(s, a, m, lda, c0.copy(), ldc, alpha, beta)
    assert_bits_equal(c_orc, c_ref, "AddMatMat")
    row_ptr, col, code, _ = O.stream_to_csr(s)
    c_csr = O.addmatmat_csr(row_ptr, col, code, s.table, s.T, n, k, a, m, lda, c0.copy(), ldc, alpha, beta)
    assert_bits_equal(c_csr, c_ref, "AddMatMat CSR restatement")
    # the reference's fp32 result sits at ~1e-7 backward error from the fp64 statement
    y64, bd = O.addmatmat_f64(s, a, m, lda, c0, ldc, alpha, beta)
    back, norm, _ = O.error_measures(c_ref, m, n, ldc, y64, bd)
    assert back < 2e-6 and norm < 1e-5
    y64b, bdb = O.addmatmat_csr_f64(row_ptr, col, code, s.table, s.T, n, k, a, m, lda, c0, ldc, alpha, beta)
    assert np.allclose(y64, y64b, rtol=1e-12, atol=1e-12) and np.allclose(bd, bdb, rtol=1e-12, atol=1e-12)


@needs_ref
@pytest.mark.parametrize("which", [0, 1, 2, 3])
def test_all_four_reference_panel_formulations_agree_with_oracle(which):
    # kernel.cc:213-369 and :771-800 compute the same panel update in different loop orders
    n, k, m, T = 256, 200, 8, 255
    codes = gen.codes_bernoulli(n, k, 0.2, seed=77)
    s = O.copyform(codes, n, k, k, gen.codebook(T, 1), T, True)
    assert s.nblocks == 1
    a = gen.dense_uniform((m, k), 2); c0 = gen.dense_uniform((m, n), 3)
    want = O.addmatmat(s, a, m, k, c0.copy(), n, 1.3, 1.0)
    if which in (2, 3):     # transposed-operand formulations work on A^T, C^T (sparse-matrix.cc:181-189)
        at = np.ascontiguousarray(a.T); ct = np.ascontiguousarray(c0.T)
        O.ref_panel_kernel(which, m, n, k, at, m, ct, m, 1.3, s.pos, s.code, s.table, T)
        got = np.ascontiguousarray(ct.T)
    else:
        got = O.ref_panel_kernel(which, m, n, k, a.copy(), k, c0.copy(), n, 1.3, s.pos, s.code, s.table, T)
    assert_bits_equal(got, want, f"panel kernel {which}")


@needs_ref
def test_transpose_kernel_shape_of_kernel_test():
    # kernel_test.cc:33-45: 1023 x 511, lda 512, ldsa 1024, bit-exact
    a = gen.dense_uniform((1024, 512), 9)
    assert_bits_equal(O.ref_trans_kernel(a, 1023, 511, 512, 1024), O.transpose(a, 1023, 511, 512, 1024), "transpose")
