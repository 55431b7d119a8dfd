"""Parity tests proper: the CUDA path, called through the C ABI, against the oracle.

Structure (upload -> CSR, decode) must be bit-exact.  Values: fp32, backward error
  max |y - y64| / (|beta c| + |alpha| sum |a||s|) <= 1e-5   (north_star tolerance)
with the reference arithmetic's own error computed alongside.
"""
import ctypes as C

import numpy as np
import pytest

import sparsematrix_b200 as sb
from sparsematrix_b200 import _capi, gen
from oracle import oracle as O
from helpers import assert_bits_equal, golden_names, golden_stream, load_golden

pytestmark = pytest.mark.gpu

TOL = 1e-5      # backward error, fp32 vs fp64 statement


def check_values(got, s, a, m, lda, c0, ldc, alpha, beta, what=""):
    n = s.cols
    y64, bound = O.addmatmat_f64(s, a, m, lda, c0, ldc, alpha, beta)
    back, norm, _ = O.error_measures(got, m, n, ldc, y64, bound)
    want = O.addmatmat(s, a, m, lda, c0.copy(), ldc, alpha, beta)
    rback, rnorm, _ = O.error_measures(want, m, n, ldc, y64, bound)
    assert back <= TOL, f"{what}: backward error {back:.3e} > {TOL} (reference arithmetic: {rback:.3e})"
    assert norm <= 1e-4, f"{what}: normwise error {norm:.3e} (reference {rnorm:.3e})"
    # padding columns of C (ldc > n) must be untouched
    if ldc > n:
        assert_bits_equal(got.reshape(m, ldc)[:, n:], c0.reshape(m, ldc)[:, n:], what + " ldc padding")
    return back, rback


def upload(s, **kw):
    return sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols, **kw)


# ------------------------------------------------------------------ golden dumps of the reference
@pytest.mark.parametrize("name", golden_names())
def test_golden_structure_and_values(name):
    g = load_golden(name)
    s = golden_stream(g)
    dev = upload(s)
    inf = dev.info()
    assert (inf.s_rows, inf.s_cols, inf.table_size) == (s.rows, s.cols, s.T)
    assert inf.fillers == int((s.code >= s.T).sum()) and inf.nnz + inf.fillers == s.entries
    row_ptr, col, code = dev.export_csr()
    o_row_ptr, o_col, o_code, _ = O.stream_to_csr(s)
    assert np.array_equal(row_ptr, o_row_ptr) and np.array_equal(col, o_col) and np.array_equal(code, o_code)
    assert_bits_equal(dev.decode_dense(int(g["stride_notrans"]), False), g["dense_notrans"], "decode NoTrans")
    assert_bits_equal(dev.decode_dense(int(g["stride_trans"]), True), g["dense_trans"], "decode Trans")
    m, lda, ldc = int(g["m"]), int(g["lda"]), int(g["ldc"])
    alpha, beta = float(g["alpha"]), float(g["beta"])
    got = dev.addmatmat_host(g["a"], m, lda, g["c_in"].copy(), ldc, alpha, beta)
    check_values(got, s, g["a"], m, lda, g["c_in"], ldc, alpha, beta, name)
    # the reference's own result is inside the same tolerance of ours
    y64, bound = O.addmatmat_f64(s, g["a"], m, lda, g["c_in"], ldc, alpha, beta)
    assert O.error_measures(g["c_out"], m, s.cols, ldc, y64, bound)[0] <= TOL
    if alpha == 0.0:        # scale only: same single rounding as the reference -> identical bits
        assert_bits_equal(got, g["c_out"], "alpha == 0")


def test_kat_selftest_through_device_path():
    # sparse-matrix.cc:209-313 re-stated in the host class and run on the GPU
    assert sb.lib().sblas_host_selftest() == 1
    g = load_golden("kat1")
    S = sb.SparseMatrix(g["codes"], 3, 2, 2, g["vals"], 8, 0)
    c = S.AddMatMat(g["a"], 1, 3, g["c_in"].copy(), 2, 1.3, 2.0)
    assert abs(c[0] - 92.513) <= 1e-3 and abs(c[1] - 44.6) <= 1e-3


# ------------------------------------------------------------------ seeded sweeps
SHAPES = [
    # n, k, density
    (257, 300, 0.10),
    (1023, 2047, 0.25),      # the harness's default shape (blas_test.cc:32)
    (1024, 512, 0.01),       # fillers
    (512, 4096, 0.001),      # filler dominated
    (300, 64, 1.0),          # fully dense
    (4096, 4096, 0.01),      # BASELINE config 1 (m = 1 below)
]
MS = [1, 2, 3, 7, 8, 9, 31, 117, 128]
SCALARS = [(1.0, 1.0), (1.0, 0.0), (1.3, 2.0), (1.8, 1.2), (0.0, 0.5)]


@pytest.mark.parametrize("shape", SHAPES)
def test_structure_bit_exact(shape):
    n, k, dens = shape
    codes = gen.codes_bernoulli(n, k, dens, seed=n * 7 + k)
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    dev = S.device_matrix()
    row_ptr, col, code = dev.export_csr()
    o_row_ptr, o_col, o_code, fill = O.stream_to_csr(s)
    assert dev.info().fillers == fill
    assert np.array_equal(row_ptr, o_row_ptr) and np.array_equal(col, o_col) and np.array_equal(code, o_code)
    assert_bits_equal(dev.decode_dense(n + 3, False), O.copyto(s, n + 3, False), "decode NoTrans")
    assert_bits_equal(dev.decode_dense(k, True), O.copyto(s, k, True), "decode Trans")


@pytest.mark.parametrize("shape", SHAPES[:5])
def test_values_all_m_and_scalars(shape):
    n, k, dens = shape
    codes = gen.codes_bernoulli(n, k, dens, seed=n * 7 + k)
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    worst = 0.0
    for mi, m in enumerate(MS):
        alpha, beta = SCALARS[mi % len(SCALARS)]
        lda, ldc = k + (mi % 3), n + (mi % 2) * 5
        a = gen.dense_uniform((m, lda), 2 + mi)
        c0 = gen.dense_uniform((m, ldc), 3 + mi)
        got = S.AddMatMat(a, m, lda, c0.copy(), ldc, alpha, beta)
        back, _ = check_values(got, s, a, m, lda, c0, ldc, alpha, beta, f"{shape} m={m} a={alpha} b={beta}")
        worst = max(worst, back)
    assert worst <= TOL


@pytest.mark.parametrize("scal", SCALARS)
@pytest.mark.parametrize("m", [1, 8, 128])
def test_values_all_scalars(m, scal):
    alpha, beta = scal
    n, k = 700, 333
    codes = gen.codes_bernoulli(n, k, 0.05, seed=41)
    codes[256:512, :] = 255                              # an unrecorded panel: those columns only get the scale
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    a = gen.dense_uniform((m, k), 42); c0 = gen.dense_uniform((m, n), 43)
    c0[0, 300] = -0.0
    got = S.AddMatMat(a, m, k, c0.copy(), n, alpha, beta)
    check_values(got, s, a, m, k, c0, n, alpha, beta, f"m={m} a={alpha} b={beta}")
    want = O.addmatmat(s, a, m, k, c0.copy(), n, alpha, beta)
    # columns of the empty panel: one rounding at most, identical to the reference bit for bit
    assert_bits_equal(got.reshape(m, n)[:, 256:512], want.reshape(m, n)[:, 256:512], "empty panel columns")


def test_config1_spmv_4096_one_percent():
    # BASELINE.json configs[0]: random 4096 x 4096, 1 % density, m = 1
    n = k = 4096
    codes = gen.codes_bernoulli(n, k, 0.01, seed=11)
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    a = gen.dense_uniform((1, k), 2); c0 = gen.dense_uniform((1, n), 3)
    got = S.AddMatMat(a, 1, k, c0.copy(), n, 1.0, 1.0)
    back, rback = check_values(got, s, a, 1, k, c0, n, 1.0, 1.0, "config 1")
    print(f"config1 backward error: gpu {back:.2e} reference arithmetic {rback:.2e}")


def test_config2_spmm_4096_ninety_percent_sparse_m128():
    # BASELINE.json configs[1]: pruned weight 4096 x 4096 at 90 % sparsity x dense 4096 x 128
    n = k = 4096; m = 128
    codes = gen.codes_bernoulli(n, k, 0.10, seed=12)
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    a = gen.dense_uniform((m, k), 2); c0 = gen.dense_uniform((m, n), 3)
    got = S.AddMatMat(a, m, k, c0.copy(), n, 1.0, 1.0)
    back, rback = check_values(got, s, a, m, k, c0, n, 1.0, 1.0, "config 2")
    print(f"config2 backward error: gpu {back:.2e} reference arithmetic {rback:.2e}")


def test_nan_in_c_propagates_with_beta_zero():
    # the reference multiplies by beta even when it is 0 (sparse-matrix.cc:149-151): NaN stays NaN
    n, k = 300, 100
    codes = gen.codes_bernoulli(n, k, 0.2, seed=9)
    S = sb.SparseMatrix(codes, n, k, k, gen.codebook(), 255, 1)
    for m in (1, 5):
        a = gen.dense_uniform((m, k), 1); c0 = gen.dense_uniform((m, n), 2)
        c0[0, 7] = np.nan; c0[m - 1, 290] = np.inf
        got = S.AddMatMat(a, m, k, c0.copy(), n, 1.0, 0.0).reshape(m, n)
        assert np.isnan(got[0, 7]) and np.isnan(got[m - 1, 290])
        assert np.isfinite(np.delete(got.reshape(-1), [7, (m - 1) * n + 290])).all()


# ------------------------------------------------------------------ direct CSR (shapes the panel format cannot hold)
def test_csr_upload_matches_stream_upload():
    n, k = 1000, 777
    codes = gen.codes_bernoulli(n, k, 0.03, seed=5)
    tab = gen.codebook()
    s = O.copyform(codes, n, k, k, tab, 255, True)
    row_ptr, col, code, _ = O.stream_to_csr(s)
    d1 = upload(s)
    d2 = sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k)
    for m in (1, 6):
        a = gen.dense_uniform((m, k), 1); c0 = gen.dense_uniform((m, n), 2)
        y1 = d1.addmatmat_host(a, m, k, c0.copy(), n, 1.3, 2.0)
        y2 = d2.addmatmat_host(a, m, k, c0.copy(), n, 1.3, 2.0)
        assert_bits_equal(y1, y2, "stream upload vs CSR upload")


def test_powerlaw_spmv_with_long_rows():
    # scaled-down BASELINE config 4: skewed rows incl. rows far above the split-row threshold
    n = k = 1 << 17
    row_ptr, col, code = gen.powerlaw_csr(n, k, seed=14, x_min=8, cap=1 << 16)
    tab = gen.codebook()
    dev = sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k)
    inf = dev.info()
    assert inf.long_rows > 0 and inf.max_row_nnz > 8192
    a = gen.dense_uniform((1, k), 1); c0 = gen.dense_uniform((1, n), 2)
    for alpha, beta in ((1.0, 1.0), (1.3, 2.0)):
        got = dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta)
        y64, bound = O.addmatmat_csr_f64(row_ptr, col, code, tab, 255, n, k, a, 1, k, c0, n, alpha, beta)
        back, norm, _ = O.error_measures(got, 1, n, n, y64, bound)
        want = O.addmatmat_csr(row_ptr, col, code, tab, 255, n, k, a, 1, k, c0.copy(), n, alpha, beta)
        rback = O.error_measures(want, 1, n, n, y64, bound)[0]
        assert back <= TOL, (back, rback)
        got2 = dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta)
        assert_bits_equal(got, got2, "run-to-run determinism")
    # SpMM on the same skewed matrix
    a = gen.dense_uniform((4, k), 1); c0 = gen.dense_uniform((4, n), 2)
    got = dev.addmatmat_host(a, 4, k, c0.copy(), n, 1.0, 1.0)
    y64, bound = O.addmatmat_csr_f64(row_ptr, col, code, tab, 255, n, k, a, 4, k, c0, n, 1.0, 1.0)
    assert O.error_measures(got, 4, n, n, y64, bound)[0] <= TOL


def test_bad_csr_and_bad_stream_are_rejected():
    tab = gen.codebook()
    row_ptr = np.array([0, 2, 3], np.int32); col = np.array([0, 9, 1], np.int32); code = np.array([1, 2, 3], np.uint8)
    with pytest.raises(sb.SblasError) as ei:
        sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, 2, 5)       # column 9 >= k
    assert ei.value.code == _capi.EFORMAT
    # a stream whose positions run past the matrix
    pos = np.array([200, 255, 255], np.uint8); cd = np.array([0, 1, 2], np.uint8)
    with pytest.raises(sb.SblasError):
        sb.DeviceMatrix.upload(pos, cd, tab, 255, [[0, 0]], [[0, 3]], 2, 256)    # rows = 2 but pos reaches row 2
    with pytest.raises(sb.SblasError):
        d = sb.DeviceMatrix.upload(np.array([1], np.uint8), np.array([0], np.uint8), tab, 255, [[0, 0]], [[0, 1]], 4, 8)
        d.addmatmat_host(np.zeros(8, np.float32), 2, 2, np.zeros(16, np.float32), 8, 1.0, 1.0)   # lda < k with m > 1
    # a single dense row never uses lda / ldc (the reference ignores them for m == 1): accepted
    d = sb.DeviceMatrix.upload(np.array([1], np.uint8), np.array([0], np.uint8), tab, 255, [[0, 0]], [[0, 1]], 4, 8)
    y = d.addmatmat_host(np.ones(4, np.float32), 1, 0, np.zeros(8, np.float32), 0, 1.0, 1.0)
    assert y[1] == tab[0] and not y[[0, 2, 3, 4, 5, 6, 7]].any()
    with pytest.raises(TypeError):
        d.addmatmat_host(np.ones(4, np.float64), 1, 4, np.zeros(8, np.float32), 8, 1.0, 1.0)      # dtype is checked


# ------------------------------------------------------------------ device-pointer entry point + slabs
def test_device_pointer_entry_matches_host_entry():
    import torch
    n, k, m = 900, 600, 33
    codes = gen.codes_bernoulli(n, k, 0.07, seed=8)
    S = sb.SparseMatrix(codes, n, k, k, gen.codebook(), 255, 1)
    dev = S.device_matrix()
    a = gen.dense_uniform((m, k + 4), 1); c0 = gen.dense_uniform((m, n + 8), 2)
    want = dev.addmatmat_host(a, m, k + 4, c0.copy(), n + 8, 1.3, 2.0)
    d_a = torch.from_numpy(a).cuda(); d_c = torch.from_numpy(c0.copy()).cuda()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        dev.addmatmat_dev(d_a, m, k + 4, d_c, n + 8, 1.3, 2.0, stream=st.cuda_stream)
    st.synchronize()
    assert_bits_equal(d_c.cpu().numpy(), want, "dev vs host entry point")


def test_output_in_registered_peer_range_gives_the_same_bits():
    """sblas_b200_peer_register: C declared to lie in peer memory takes the epilogues for NVLink-mapped output
    (no L2 prefetch of C; for beta == 1 the sums are added by red.global.add.f32, C never read) -- every SpMV
    kernel and the SpMM tile kernel, against the same product into an undeclared buffer.  On one GPU the "peer"
    range is local memory, which the reductions serve just the same (two GPUs: tests/test_peer_output_gpu.py)."""
    import torch
    L = sb.lib()
    tab = gen.codebook()
    mats = []
    codes = gen.codes_bernoulli(3000, 1500, 0.06, seed=21)
    codes[256:512] = 255                                                     # an empty panel
    mats.append(("rows", sb.SparseMatrix(codes, 3000, 1500, 1500, tab, 255, 1).device_matrix(), 3000, 1500))
    # skewed rows, the length-sorted copy forced, split long rows, rows without entries, n not a multiple of 4: its
    # row sums are formed in a local buffer and added out in row order (run_spmv_sorted_to_peer)
    n, k = (1 << 16) + 7, 1 << 16
    row_ptr, col, code = gen.powerlaw_csr(n, k, seed=14, x_min=8, cap=1 << 15)
    lens = np.diff(row_ptr).astype(np.int64)
    keep = np.ones(col.shape[0], bool)
    keep[row_ptr[10]:row_ptr[17]] = False; keep[row_ptr[n - 2]:row_ptr[n - 1]] = False
    lens[10:17] = 0; lens[n - 2] = 0
    col, code = col[keep], code[keep]
    row_ptr = np.concatenate(([0], np.cumsum(lens))).astype(np.int32)
    skew = _with_env("SBLAS_B200_SORT_ROWS", "1", lambda: sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k))
    assert skew.info().long_rows > 0
    mats.append(("sorted+long", skew, n, k))
    codes = gen.codes_bernoulli(1100, 2048, 0.5, seed=22)
    mats.append(("dense codes", sb.SparseMatrix(codes, 1100, 2048, 2048, tab, 255, 1).device_matrix(), 1100, 2048))
    for name, dev, n, k in mats:
        for m in (1, 8, 64):
            if name == "sorted+long" and m == 64:
                continue
            a = torch.from_numpy(gen.dense_uniform((m, k), 31 + m)).cuda()
            c0 = torch.from_numpy(gen.dense_uniform((m, n), 32 + m)).cuda()
            for alpha, beta in ((1.0, 1.0), (0.7, 1.0), (1.3, 2.0)):
                want = c0.clone()
                dev.addmatmat_dev(a, m, k, want, n, alpha, beta)
                for shift in ((0, 1) if m == 1 else (0,)):                      # shift 1: y off the 16-byte grid
                    buf = torch.empty(m * n + shift, device="cuda")
                    got = buf[shift:].view(m, n)
                    got.copy_(c0)
                    assert L.sblas_b200_peer_register(C.c_void_p(buf.data_ptr()), buf.numel() * 4) == _capi.OK
                    try:
                        dev.addmatmat_dev(a, m, k, got, n, alpha, beta)
                        torch.cuda.synchronize()
                    finally:
                        assert L.sblas_b200_peer_unregister(C.c_void_p(buf.data_ptr())) == _capi.OK
                    assert_bits_equal(got.cpu().numpy(), want.cpu().numpy(),
                                      f"{name} m={m} alpha={alpha} beta={beta} shift={shift}")


@pytest.mark.parametrize("parts", [2, 3, 8])
def test_virtual_shards_concatenate_to_unsharded_result(parts):
    # run the slabs one after another on one device; every output row is computed
    # by its own slab alone, so the concatenation equals the unsharded result (bit-identical for SpMM)
    n, k = 2048 + 100, 500
    codes = gen.codes_bernoulli(n, k, 0.05, seed=77)
    S = sb.SparseMatrix(codes, n, k, k, gen.codebook(), 255, 1)
    _, _, _, _, bb, bib = S.members()
    splits = sb.partition(bb, bib, n, parts)
    for m in (1, 16):
        a = gen.dense_uniform((m, k), 1); c0 = gen.dense_uniform((m, n), 2)
        whole = S.AddMatMat(a, m, k, c0.copy(), n, 1.3, 2.0)
        pieces = c0.copy()
        tot_nnz = 0
        for g in range(parts):
            lo, hi = int(splits[g]), int(splits[g + 1])
            if hi == lo:
                continue
            slab = S.upload_slab(lo, hi)
            tot_nnz += slab.info().nnz
            view = pieces.reshape(m, n)[:, lo:]                 # c points at column col_begin, ldc stays n
            ptr = view.ctypes.data
            _capi.check(sb.lib().sblas_b200_addmatmat_host(slab._h, a.ctypes.data_as(C.c_void_p), m, k,
                                                           C.c_void_p(ptr), n, 1.3, 2.0))
        assert tot_nnz == S.device_matrix().info().nnz
        if m == 1:
            # SpMV reads aligned 4-entry blocks, so which lane sums which entry depends on where a
            # row starts in the slab's own arrays: same values to rounding, not to the bit
            s = O.copyform(codes, n, k, k, gen.codebook(), 255, True)
            check_values(pieces, s, a, m, k, c0, n, 1.3, 2.0, f"virtual shards parts={parts} m=1")
            assert np.max(np.abs(pieces - whole)) <= 1e-5 * np.max(np.abs(whole))
        else:
            assert_bits_equal(pieces, whole, f"virtual shards parts={parts} m={m}")


# ------------------------------------------------------------------ the C++ harness (src/test), same library
@pytest.mark.parametrize("binary,expect", [("sparse-matrix_test", "SparseMatrix Test success"),
                                           ("kernel_test", "Kernel Test success"),
                                           ("parity_test", "Parity Test success")])
def test_cpp_harness_binaries(binary, expect):
    import os
    import subprocess
    path = os.path.join(os.path.dirname(sb.LIB_PATH), "..", "bin", binary)
    if not os.path.exists(path):
        pytest.skip(f"{binary} not built")
    out = subprocess.run([path], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and expect in out.stdout, out.stdout[-2000:] + out.stderr[-500:]


# ------------------------------------------------------------------ the reference's free kernel functions (kernel.h:42-62)
def test_free_kernel_functions_against_the_reference_library():
    """sblas_trans_kernel / sblas_beta_operation_kernel / the four panel formulations, device backed
    (sblas_b200_transpose_host, _scale_host, _panel_product_host), against the unmodified reference's own
    functions on the same inputs: transposes and the scale bit for bit, panel products within the contract."""
    L = sb.lib()
    # kernel_test.cc:33-45: 1023 x 511, lda 512, ldsa 1024
    a = gen.dense_uniform((1024, 512), 9)
    sa = np.full((511, 1024), -3.0, np.float32)
    _capi.check(L.sblas_b200_transpose_host(a.ctypes.data_as(C.c_void_p), 1023, 511, 512, sa.ctypes.data_as(C.c_void_p), 1024))
    want = np.full((511, 1024), -3.0, np.float32); want[:, :1023] = a[:1023, :511].T
    assert_bits_equal(sa, want, "transpose (untouched padding included)")
    if O.have_ref():
        assert_bits_equal(sa[:, :1023], O.ref_trans_kernel(a, 1023, 511, 512, 1024)[:, :1023], "transpose vs reference")
    c = gen.dense_uniform((37, 320), 10); c_want = c.copy(); c_want[:, :301] *= np.float32(1.7)
    _capi.check(L.sblas_b200_scale_host(c.ctypes.data_as(C.c_void_p), 37, 301, 320, 1.7))
    assert_bits_equal(c, c_want, "beta scale")
    # one panel with fillers (2 % density), every formulation
    n, k, m, T = 256, 300, 11, 255
    codes = gen.codes_bernoulli(n, k, 0.02, seed=78)
    tab = gen.codebook(T, 1)
    s = O.copyform(codes, n, k, k, tab, T, True)
    assert s.nblocks == 1 and int((s.code >= T).sum()) > 0
    am = gen.dense_uniform((m, k), 2); c0 = gen.dense_uniform((m, n), 3)
    y64, bound = O.addmatmat_f64(s, am, m, k, c0, n, 1.3, 1.0)
    for trans in (0, 1):
        if trans:
            at = np.zeros((k, 16), np.float32); at[:, :m] = am.T
            ct = np.zeros((n, 16), np.float32); ct[:, :m] = c0.T
            _capi.check(L.sblas_b200_panel_product_host(1, m, n, k, at.ctypes.data_as(C.c_void_p), 16, ct.ctypes.data_as(C.c_void_p), 16,
                                                        1.3, s.pos.ctypes.data_as(C.c_void_p), s.code.ctypes.data_as(C.c_void_p),
                                                        int(s.pos.shape[0]), s.table.ctypes.data_as(C.c_void_p), T))
            got = np.ascontiguousarray(ct[:, :m].T)
            assert np.all(ct[:, m:] == 0.0)                     # the padding of aligned_m stays untouched
        else:
            got = c0.copy()
            _capi.check(L.sblas_b200_panel_product_host(0, m, n, k, am.ctypes.data_as(C.c_void_p), k, got.ctypes.data_as(C.c_void_p), n,
                                                        1.3, s.pos.ctypes.data_as(C.c_void_p), s.code.ctypes.data_as(C.c_void_p),
                                                        int(s.pos.shape[0]), s.table.ctypes.data_as(C.c_void_p), T))
        assert O.error_measures(got, m, n, n, y64, bound)[0] <= TOL, trans
        if O.have_ref():
            ref = O.ref_panel_kernel(1, m, n, k, am.copy(), k, c0.copy(), n, 1.3, s.pos, s.code, s.table, T)
            assert np.max(np.abs(ref - got)) <= 1e-5 * float(np.max(np.abs(ref)))


# ------------------------------------------------------------------ host entry point, buffers of the library's allocator
def test_host_alloc_buffers_take_the_overlapped_route_with_the_same_bits():
    n, k, m = 1500, 900, 33
    codes = gen.codes_bernoulli(n, k, 0.07, seed=41)
    S = sb.SparseMatrix(codes, n, k, k, gen.codebook(), 255, 1)
    a = gen.dense_uniform((m, k), 1); c0 = gen.dense_uniform((m, n), 2)
    want = S.AddMatMat(a, m, k, c0.copy(), n, 1.3, 0.7)                 # pageable numpy buffers
    ha = sb.host_buffer((m, k)); hc = sb.host_buffer((m, n))
    ha[...] = a; hc[...] = c0
    S.AddMatMat(ha, m, k, hc, n, 1.3, 0.7)
    assert_bits_equal(hc, want, "sblas_b200_host_alloc buffers vs pageable")
    # m = 1 with small operands: x staged by a kernel through the mapped pointer, y updated in place (no DMA)
    for alpha, beta in ((1.0, 1.0), (1.3, 0.7), (0.0, 2.0)):
        want1 = S.AddMatMat(a[:1].copy(), 1, k, c0[:1].copy(), n, alpha, beta)
        hx = sb.host_buffer((1, k)); hy = sb.host_buffer((1, n))
        hx[...] = a[:1]; hy[...] = c0[:1]
        S.AddMatMat(hx, 1, k, hy, n, alpha, beta)
        assert_bits_equal(hy, want1, f"direct small SpMV route alpha={alpha} beta={beta}")
    hc2 = sb.host_buffer((m, n + 8))                                    # ldc > n: the padding stays untouched
    hc2[...] = 7.0; hc2[:, :n] = c0
    S.AddMatMat(ha, m, k, hc2, n + 8, 1.3, 0.7)
    assert_bits_equal(hc2[:, :n], want, "host_alloc buffer with ldc > n")
    assert np.all(hc2[:, n:] == 7.0)


# ------------------------------------------------------------------ host entry point, pinned caller buffers
def test_pinned_host_buffers_give_the_same_bits_as_pageable_ones():
    """Pinned A/C take the overlapped route (product into a side buffer while C travels in, one
    combine kernel writing into the caller's C); pageable buffers take copy-product-copy.  Both
    must produce identical bits, incl. columns without entries, -0, ragged ldc and repeated calls."""
    import torch
    n, k = 777, 1000
    codes = gen.codes_bernoulli(n, k, 0.04, seed=21)
    codes[5, :] = 255; codes[300:320, :] = 255                 # S columns without entries
    tab = gen.codebook()
    dev = upload(O.copyform(codes, n, k, k, tab, 255, True))
    for m, lda, ldc in ((1, k, n), (3, k + 4, n + 3), (40, k, n + 4), (130, k + 8, n)):
        a = gen.dense_uniform((m, lda), 3 + m); c0 = gen.dense_uniform((m, ldc), 4 + m)
        c0[0, 5] = -0.0; c0[0, 6] = -0.0
        for alpha, beta in ((1.0, 1.0), (1.3, 2.0), (0.7, 0.0)):
            want = dev.addmatmat_host(a, m, lda, c0.copy(), ldc, alpha, beta)
            for rep in range(2):                                # second call: side buffer reused
                pa = torch.from_numpy(a.copy()).pin_memory(); pc = torch.from_numpy(c0.copy()).pin_memory()
                dev.addmatmat_host(pa, m, lda, pc, ldc, alpha, beta)
                assert_bits_equal(pc.numpy(), want, f"pinned vs pageable m={m} alpha={alpha} beta={beta} rep={rep}")
    # k large enough that m = 300 is cut into slices (copies in / products / copies out pipelined)
    n2, k2 = 520, 9000
    codes2 = gen.codes_bernoulli(n2, k2, 0.02, seed=22)
    dev2 = upload(O.copyform(codes2, n2, k2, k2, tab, 255, True))
    m = 300
    a = gen.dense_uniform((m, k2), 5); c0 = gen.dense_uniform((m, n2 + 4), 6)
    want = dev2.addmatmat_host(a, m, k2, c0.copy(), n2 + 4, 1.3, 2.0)
    pa = torch.from_numpy(a).pin_memory(); pc = torch.from_numpy(c0.copy()).pin_memory()
    dev2.addmatmat_host(pa, m, k2, pc, n2 + 4, 1.3, 2.0)
    assert_bits_equal(pc.numpy(), want, "sliced pinned path vs pageable")
    # unaligned pinned C (scalar combine) after a larger call (side buffer larger than needed)
    m = 5
    a = gen.dense_uniform((m, k), 9); c0 = gen.dense_uniform((m * n + 1,), 10)
    want = dev.addmatmat_host(a, m, k, c0[1:].copy(), n, 1.0, 1.0)
    pc = torch.from_numpy(c0.copy()).pin_memory()
    dev.addmatmat_host(torch.from_numpy(a).pin_memory(), m, k, pc[1:], n, 1.0, 1.0)
    assert_bits_equal(pc[1:].numpy(), want, "unaligned pinned C")
    assert pc[0].item() == c0[0]


# ------------------------------------------------------------------ CopyForm on the device
def _assert_same_streams(pos, code, bb, bib, s, what):
    assert pos.shape[0] == s.entries, (what, pos.shape[0], s.entries)
    assert np.array_equal(pos, s.pos), what + ": pos_index"
    assert np.array_equal(code, s.code), what + ": val_index"
    assert np.array_equal(bb, s.block_bounds.reshape(-1, 2)), what + ": block_bounds"
    assert np.array_equal(bib, s.block_index_bounds.reshape(-1, 2)), what + ": block_index_bounds"


@pytest.mark.parametrize("rows,cols,stride,dens,T,trans", [
    (3, 2, 2, 0.5, 8, False),                  # SelfTest size
    (1023, 511, 512, 0.25, 63, False),         # SelfTest round trip shape, narrow last panel
    (1023, 511, 512, 0.25, 63, True),
    (777, 1000, 1003, 0.04, 255, True),        # odd stride, several panels
    (600, 2000, 2000, 0.001, 255, False),      # hypersparse: long filler runs, empty panels
    (600, 2000, 2000, 0.001, 255, True),
    (257, 513, 520, 1.0, 200, False),          # fully dense, codes >= T count as zeros
    (4096, 300, 300, 0.10, 255, True),
    (40, 70000, 70000, 0.0005, 17, False),     # many panels, most of them empty
])
def test_device_copyform_streams_are_bit_identical(rows, cols, stride, dens, T, trans):
    rng = np.random.default_rng(rows * 31 + cols)
    codes = np.full((rows, stride), 255, np.uint8)
    mask = rng.random((rows, cols)) < dens
    vals = rng.integers(0, 256 if dens == 1.0 else T, size=(rows, cols)).astype(np.uint8)
    codes[:, :cols][mask] = vals[mask]
    tab = gen.codebook(T, 1)
    s = O.copyform(codes, rows, cols, stride, tab, T, trans)
    pos, code, bb, bib, dev = sb.copyform_device(codes, rows, cols, stride, tab, T, trans)
    _assert_same_streams(pos, code, bb, bib, s, "device CopyForm vs oracle")
    # the device matrix built from the device-resident streams = the one uploaded from host streams
    ref = upload(s)
    for x, y in zip(dev.export_csr(), ref.export_csr()):
        assert np.array_equal(x, y)
    assert dev.info().fillers == ref.info().fillers


def test_device_copyform_edge_cases_and_host_class_route():
    import os
    tab = gen.codebook(255, 1)
    # nothing valid at all -> no entries, no panels; T = 0 -> empty object
    codes = np.full((50, 300), 255, np.uint8)
    pos, code, bb, bib, dev = sb.copyform_device(codes, 50, 300, 300, tab, 255, False)
    assert pos.shape[0] == 0 and bb.shape[0] == 0 and dev.nnz == 0
    pos, code, bb, bib, dev = sb.copyform_device(codes, 50, 300, 300, tab, 0, True)
    assert pos.shape[0] == 0 and bb.shape[0] == 0
    # single element at the far corner: one long filler run
    codes[49, 299] = 3
    s = O.copyform(codes, 50, 300, 300, tab, 255, False)
    pos, code, bb, bib, dev = sb.copyform_device(codes, 50, 300, 300, tab, 255, False)
    _assert_same_streams(pos, code, bb, bib, s, "corner element")
    # the host class takes the device route for large inputs (and when forced) and ends up equal to the host scan
    n, k = 700, 900
    big = gen.codes_bernoulli(n, k, 0.05, seed=31)
    a = gen.dense_uniform((5, k), 1); c0 = gen.dense_uniform((5, n), 2)
    old = os.environ.get("SBLAS_B200_COPYFORM")
    try:
        os.environ["SBLAS_B200_COPYFORM"] = "host"
        H = sb.SparseMatrix(big, n, k, k, tab, 255, sb.SparseMatrix.Trans)
        os.environ["SBLAS_B200_COPYFORM"] = "device"
        D = sb.SparseMatrix(big, n, k, k, tab, 255, sb.SparseMatrix.Trans)
    finally:
        if old is None: os.environ.pop("SBLAS_B200_COPYFORM", None)
        else: os.environ["SBLAS_B200_COPYFORM"] = old
    assert H == D
    for x, y in zip(H.members(), D.members()):
        assert np.array_equal(np.asarray(x), np.asarray(y))
    assert_bits_equal(D.AddMatMat(a, 5, k, c0.copy(), n, 1.3, 2.0), H.AddMatMat(a, 5, k, c0.copy(), n, 1.3, 2.0), "device-built vs host-built object")
    assert np.array_equal(D.CopyTo(k, sb.SparseMatrix.Trans), H.CopyTo(k, sb.SparseMatrix.Trans))    # S^T is n x k: stride >= k


def test_pageable_buffers_seen_twice_are_registered_and_give_the_same_bits():
    """The registration cache of the host entry point (pinned_memory.cu): the first call with malloc'ed buffers is staged,
    the second registers them, later ones take the overlapped route -- every call must produce the first call's bits."""
    import os
    n, k, m = 2048, 1024, 96                        # a: 384 KiB, c: 768 KiB (both above the 64 KiB floor)
    old_env = os.environ.get("SBLAS_B200_HOST_REGISTER")
    os.environ["SBLAS_B200_HOST_REGISTER"] = "1"     # opt-in, read per call
    try:
        _registered_pageable_case(n, k, m)
    finally:
        sb.lib().sblas_b200_host_forget(None)
        if old_env is None: os.environ.pop("SBLAS_B200_HOST_REGISTER", None)
        else: os.environ["SBLAS_B200_HOST_REGISTER"] = old_env


def _registered_pageable_case(n, k, m):
    codes = gen.codes_bernoulli(n, k, 0.05, seed=51)
    tab = gen.codebook(255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    dev = upload(s)
    a = gen.dense_uniform((m, k), 52)
    c0 = gen.dense_uniform((m, n), 53)
    c = np.empty_like(c0)
    outs = []
    for call in range(4):
        c[...] = c0
        dev.addmatmat_host(a, m, k, c, n, 1.3, 2.0)
        outs.append(c.copy())
    for call in range(1, 4):
        assert_bits_equal(outs[call], outs[0], f"call {call + 1} with the same pageable buffers")
    check_values(outs[3], s, a, m, k, c0, n, 1.3, 2.0, "registered pageable buffers")
    # a strided C (ldc > n) through the same mechanism; padding untouched
    ldc = n + 8
    c2 = gen.dense_uniform((m, ldc), 54)
    c2_0 = c2.copy()
    for call in range(3):
        c2[...] = c2_0
        dev.addmatmat_host(a, m, k, c2, ldc, 1.0, 1.0)
    check_values(c2, s, a, m, k, c2_0, ldc, 1.0, 1.0, "registered strided C")
    assert sb.lib().sblas_b200_host_forget(a.ctypes.data_as(C.c_void_p)) == 0
    c[...] = c0
    dev.addmatmat_host(a, m, k, c, n, 1.3, 2.0)                     # a is pageable again, c still registered
    assert_bits_equal(c, outs[0], "after host_forget(a)")
    dev.free()                                                       # drops the remaining registrations
    assert sb.lib().sblas_b200_host_forget(None) == 0


def test_device_copyform_reads_no_byte_past_the_last_row():
    """A strided view whose last row ends at the end of its allocation: (rows-1)*stride + cols bytes, not rows*stride
    (ADVICE r01).  The buffer ends flush against a PROT_NONE page, so an over-read faults."""
    import ctypes
    import mmap
    rows, cols, stride, T = 600, 500, 512, 255                 # rows*cols >= 2^18: the host class takes the device route
    need = (rows - 1) * stride + cols
    page = mmap.PAGESIZE
    span = (need + page - 1) // page * page
    mm = mmap.mmap(-1, span + page)
    base = ctypes.addressof(ctypes.c_char.from_buffer(mm))
    libc = ctypes.CDLL(None, use_errno=True)
    assert libc.mprotect(ctypes.c_void_p(base + span), ctypes.c_size_t(page), 0) == 0        # guard page
    buf = np.frombuffer(mm, dtype=np.uint8, count=need, offset=span - need)
    full = np.full((rows, stride), 255, np.uint8)
    full[:, :cols] = gen.codes_bernoulli(rows, cols, 0.1, seed=41)
    buf[:] = full.reshape(-1)[:need]
    tab = gen.codebook(T, 1)
    want = O.copyform(full, rows, cols, stride, tab, T, False)
    pos, code, bb, bib, dev = sb.copyform_device(buf, rows, cols, stride, tab, T, False)
    _assert_same_streams(pos, code, bb, bib, want, "minimal-size strided input")
    S = sb.SparseMatrix(buf, rows, cols, stride, tab, T, sb.SparseMatrix.NoTrans)
    p2, c2, _, _, _, _ = S.members()
    assert np.array_equal(p2, want.pos) and np.array_equal(c2, want.code)
    del buf, S
    libc.mprotect(ctypes.c_void_p(base + span), ctypes.c_size_t(page), 3)


def test_decode_dense_into_device_memory():
    import torch
    n, k = 300, 500
    codes = gen.codes_bernoulli(n, k, 0.05, seed=33)
    tab = gen.codebook()
    dev = upload(O.copyform(codes, n, k, k, tab, 255, True))
    for trans, shape in ((False, (k, n + 5)), (True, (n, k + 3))):
        want = dev.decode_dense(shape[1], trans)
        d = torch.full(shape, 7.0, dtype=torch.float32, device="cuda")
        _capi.check(sb.lib().sblas_b200_decode_dense(dev._h, C.c_void_p(d.data_ptr()), shape[1], int(trans)))
        assert np.array_equal(d.cpu().numpy(), want)


# ------------------------------------------------------------------ every template instance, forced through the experiment switches
def _with_env(name, value, fn):
    import os
    old = os.environ.get(name)
    os.environ[name] = value
    try:
        return fn()
    finally:
        if old is None: os.environ.pop(name, None)
        else: os.environ[name] = old


def test_every_spmm_tile_variant_gives_the_same_bits():
    """rows per slot x float4s per lane x warps: every row is summed by one slot in column order,
    so all instances of spmm_tile_kernel must agree bit for bit, and with the fp64 statement to TOL."""
    n, k = 610, 1300                                           # 6 chunks of 255 X rows, ragged last group
    codes = gen.codes_bernoulli(n, k, 0.08, seed=41)
    codes[7, :] = 255; codes[100:109, :] = 255                 # empty W rows
    codes[300, :] = (np.arange(k) % 200).astype(np.uint8)      # one dense row
    tab = gen.codebook()
    s = O.copyform(codes, n, k, k, tab, 255, True)
    dev = upload(s)
    for m, lda, ldc, alpha, beta in ((17, k, n, 1.0, 1.0), (40, k + 4, n + 3, 1.3, 2.0), (100, k, n, 0.7, 0.0)):
        a = gen.dense_uniform((m, lda), 50 + m); c0 = gen.dense_uniform((m, ldc), 60 + m)
        y64, bound = O.addmatmat_f64(s, a, m, lda, c0, ldc, alpha, beta)
        first = None
        for variant in ("1104", "1128", "1204", "1216", "2103", "2116", "2204", "2216"):
            got = _with_env("SBLAS_B200_TILE", variant, lambda: dev.addmatmat_host(a, m, lda, c0.copy(), ldc, alpha, beta))
            back = O.error_measures(got, m, n, ldc, y64, bound)[0]
            assert back <= TOL, (variant, m, back)
            if first is None: first = got
            else: assert_bits_equal(got, first, f"tile variant {variant} vs 1104, m={m}")


def test_every_spmv_width_and_the_sorted_copy():
    n, k = 70000, 3000                                         # >= 65536 rows so the sorted copy can be forced
    row_ptr, col, code = gen.powerlaw_csr(n, k, seed=42, x_min=2, cap=2500)
    lens = np.diff(row_ptr).astype(np.int64)                   # rows 10..19 and the last one lose their entries
    keep = np.ones(col.shape[0], bool)
    keep[row_ptr[10]:row_ptr[20]] = False; keep[row_ptr[n - 1]:] = False
    lens[10:20] = 0; lens[n - 1] = 0
    col, code = col[keep], code[keep]
    row_ptr = np.concatenate(([0], np.cumsum(lens))).astype(np.int32)
    tab = gen.codebook()
    a = gen.dense_uniform((1, k), 1); c0 = gen.dense_uniform((1, n), 2)
    y64, bound = O.addmatmat_csr_f64(row_ptr, col, code, tab, 255, n, k, a, 1, k, c0, n, 1.3, 2.0)
    for sort_rows in ("0", "1"):
        for lanes in ("1", "2", "4", "8", "16", "32"):
            dev = _with_env("SBLAS_B200_SORT_ROWS", sort_rows, lambda: _with_env(
                "SBLAS_B200_SPMV_LANES", lanes, lambda: sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k)))
            assert dev.info().spmv_lanes == int(lanes)
            got = dev.addmatmat_host(a, 1, k, c0.copy(), n, 1.3, 2.0)
            back = O.error_measures(got, 1, n, n, y64, bound)[0]
            assert back <= TOL, (sort_rows, lanes, back)
            # rows without entries keep exactly beta*c
            empty = np.flatnonzero(np.diff(row_ptr) == 0)
            if empty.size:
                assert np.array_equal(got.reshape(-1)[empty], (np.float32(2.0) * c0.reshape(-1)[empty]))


def test_long_row_spmv_full_warp_per_row():
    """Mean row length >= 128 selects a full warp per row: rows of exactly 128 / 129 entries, a full-width
    row, an empty row and a single-entry last row."""
    n, k = 333, 5000
    codes = gen.codes_bernoulli(n, k, 0.30, seed=51)
    codes[3, :] = 255                                          # empty row
    codes[4, :] = 255; codes[4, :128] = 7                      # exactly one full record
    codes[5, :] = 255; codes[5, :129] = 9                      # one full record + 1 entry
    codes[6, :] = (np.arange(k) % 250).astype(np.uint8)        # every column: 40 records
    codes[n - 1, :] = 255; codes[n - 1, k - 1] = 1             # single entry in the last row
    tab = gen.codebook()
    s = O.copyform(codes, n, k, k, tab, 255, True)
    dev = upload(s)
    assert dev.info().spmv_lanes == 32
    a = gen.dense_uniform((1, k), 1); c0 = gen.dense_uniform((1, n), 2); c0[0, 3] = -0.0
    for alpha, beta in ((1.0, 1.0), (1.3, 2.0), (0.7, 0.0)):
        # (30 % of the elements are stored: without the switch this matrix takes the dense-code kernel)
        got = _with_env("SBLAS_B200_SPMV_DENSE", "0", lambda: dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta))
        y64, bound = O.addmatmat_f64(s, a, 1, k, c0, n, alpha, beta)
        back = O.error_measures(got, 1, n, n, y64, bound)[0]
        want = O.addmatmat(s, a, 1, k, c0.copy(), n, alpha, beta)
        assert back <= TOL, (alpha, beta, back)
        assert got.reshape(-1)[3].tobytes() == want.reshape(-1)[3].tobytes()       # empty row: exactly beta*c (keeps -0)
        got2 = _with_env("SBLAS_B200_SPMV_DENSE", "0", lambda: dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta))
        assert_bits_equal(got, got2, "run-to-run determinism")
    assert dev.info().device_bytes > 0


@pytest.mark.parametrize("n,k,dens", [(333, 5000, 0.30), (1000, 700, 0.5), (515, 1025, 0.25), (64, 16384, 0.2),
                                      (3000, 512, 0.9), (5, 40, 0.6), (2000, 3000, 0.01)])
def test_dense_code_spmv(n, k, dens):
    """m = 1 on matrices with >= 3/16 of their elements stored (spmv_dense.cu): one byte per element, x in
    registers.  Values against fp64, rows without entries, every column of a row, an Inf / NaN in x where a
    row has no entry (must not reach that row) and where it has one (must), determinism, and the row kernels
    on the same handle as a cross-check.  The last case is too sparse for the rule and is forced."""
    codes = gen.codes_bernoulli(n, k, dens, seed=71)
    codes[min(3, n - 1), :] = 255                              # empty row
    if n > 6:
        codes[6, :] = (np.arange(k) % 250).astype(np.uint8)    # every column stored
    codes[:, k // 2] = 255                                     # a column no entry uses ...
    codes[0, k // 2] = 9                                       # ... except row 0
    tab = gen.codebook()
    s = O.copyform(codes, n, k, k, tab, 255, True)
    dev = upload(s)
    force = "1" if dens < 0.19 else ""
    a = gen.dense_uniform((1, k), 1); c0 = gen.dense_uniform((1, n), 2); c0[0, min(3, n - 1)] = -0.0
    bytes_before = dev.info().device_bytes
    for alpha, beta in ((1.0, 1.0), (1.3, 2.0), (0.7, 0.0)):
        got = _with_env("SBLAS_B200_SPMV_DENSE", force, lambda: dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta))
        y64, bound = O.addmatmat_f64(s, a, 1, k, c0, n, alpha, beta)
        back = O.error_measures(got, 1, n, n, y64, bound)[0]
        assert back <= TOL, (alpha, beta, back)
        want = O.addmatmat(s, a, 1, k, c0.copy(), n, alpha, beta)
        e = min(3, n - 1)
        assert got.reshape(-1)[e].tobytes() == want.reshape(-1)[e].tobytes()       # empty row: exactly beta*c (keeps -0)
        again = _with_env("SBLAS_B200_SPMV_DENSE", force, lambda: dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta))
        assert_bits_equal(got, again, "run-to-run determinism")
        rows = _with_env("SBLAS_B200_SPMV_DENSE", "0", lambda: dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta))
        assert O.error_measures(rows, 1, n, n, y64, bound)[0] <= TOL
        assert np.max(np.abs(rows - got)) <= 2e-5 * max(1.0, float(np.max(np.abs(got))))
        for bad in (np.inf, np.nan):
            a_bad = a.copy(); a_bad[0, k // 2] = bad
            got_bad = _with_env("SBLAS_B200_SPMV_DENSE", force, lambda: dev.addmatmat_host(a_bad, 1, k, c0.copy(), n, alpha, beta))
            assert not np.isfinite(got_bad.reshape(-1)[0])                          # row 0 uses that column
            assert_bits_equal(got_bad.reshape(-1)[1:], got.reshape(-1)[1:], f"{bad} in a column only row 0 uses")
    assert dev.info().device_bytes >= bytes_before + n * k    # the dense code matrix was built and is accounted for


def test_spmv_never_touches_x_where_no_entry_is():
    """Every sub-warp width on a matrix with wide gaps, rows of 4 j and 4 j + 1 entries (block edges of the
    128-bit column loads), empty rows; an Inf in x at a column no entry uses must not reach any result
    (padding / masked lanes must not multiply 0 * x)."""
    n, k = 500, 3000
    rng = np.random.default_rng(61)
    codes = gen.codes_bernoulli(n, k, 0.06, seed=61)
    codes[:, 700:1400] = 255                                   # a 700-column gap in every row: 2 fillers each
    codes[10, :] = 255                                         # empty row
    codes[11, :] = 255; codes[11, 2990:2994] = 5               # 4 entries, first one after 11 fillers
    codes[12, :] = 255; codes[12, 0:5] = 6                     # 5 entries from column 0
    codes[13, :] = 255; codes[13, k - 1] = 7                   # single entry in the last column
    codes[:, 1000] = 255                                       # column 1000 is used by no entry
    tab = gen.codebook()
    s = O.copyform(codes, n, k, k, tab, 255, True)
    a = gen.dense_uniform((1, k), 1); c0 = gen.dense_uniform((1, n), 2); c0[0, 10] = -0.0
    a_inf = a.copy(); a_inf[0, 1000] = np.inf
    for lanes in ("1", "2", "4", "8", "16", "32"):
        dev = _with_env("SBLAS_B200_SPMV_LANES", lanes, lambda: upload(s))
        for alpha, beta in ((1.0, 1.0), (1.3, 2.0), (0.7, 0.0)):
            got = dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta)
            y64, bound = O.addmatmat_f64(s, a, 1, k, c0, n, alpha, beta)
            assert O.error_measures(got, 1, n, n, y64, bound)[0] <= TOL, (lanes, alpha, beta)
            want = O.addmatmat(s, a, 1, k, c0.copy(), n, alpha, beta)
            assert got.reshape(-1)[10].tobytes() == want.reshape(-1)[10].tobytes()   # empty row: exactly beta*c
            got_inf = dev.addmatmat_host(a_inf, 1, k, c0.copy(), n, alpha, beta)
            assert_bits_equal(got_inf, got, f"Inf in an unused column of x, lanes={lanes}")


# ------------------------------------------------------------------ BASELINE full sizes, checked through sampled rows and invariants
def _sampled_backward_error(got_cols, w_rows_codes, tab, a, c0_cols, alpha, beta):
    """got_cols: m x s outputs of the sampled W rows; w_rows_codes: s x k codes of those rows."""
    w = np.where(w_rows_codes < 255, tab.astype(np.float64)[np.minimum(w_rows_codes, 254)], 0.0)      # s x k
    a64 = a.astype(np.float64)
    y64 = beta * c0_cols.astype(np.float64) + alpha * (a64 @ w.T)
    bound = np.abs(beta * c0_cols.astype(np.float64)) + abs(alpha) * (np.abs(a64) @ np.abs(w).T)
    return float(np.max(np.abs(got_cols.astype(np.float64) - y64) / np.maximum(bound, 1e-300)))


def test_config3_full_size_sampled_rows_and_invariants():
    # BASELINE.json configs[2]: 16384 x 16384 at 95 % sparsity x dense N = 512
    import torch
    n = k = 16384; m = 512
    codes = gen.codes_bernoulli(n, k, 0.05, seed=13)
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, 1)           # CopyForm takes the device route at this size
    dev = S.device_matrix()
    assert dev.nnz == int((codes != 255).sum())
    a = gen.dense_uniform((m, k), 2); c0 = gen.dense_uniform((m, n), 3)
    got = dev.addmatmat_host(a, m, k, c0.copy(), n, 1.3, 2.0)
    family = dev.last_product()["kernel"].split("<")[0]
    rows = np.random.default_rng(5).choice(n, 48, replace=False)
    err = _sampled_backward_error(got[:, rows], codes[rows], tab, a, c0[:, rows], 1.3, 2.0)
    assert err <= TOL, err
    # an output element does not depend on how many other A rows are multiplied with it (m-tile
    # shape, kernel variant) nor on the host route (pinned, sliced copies) -- within one kernel family: the tensor-core
    # form (chosen here for m = 512) and the tile form (chosen for a 64-row slice) add in different orders
    part = dev.addmatmat_host(a[100:164], 64, k, c0[100:164].copy(), n, 1.3, 2.0)
    if dev.last_product()["kernel"].split("<")[0] == family:
        assert_bits_equal(part, got[100:164], "m = 64 slice vs rows 100..163 of the m = 512 product")
    else:
        assert _sampled_backward_error(part[:, rows], codes[rows], tab, a[100:164], c0[100:164][:, rows], 1.3, 2.0) <= TOL
    part = dev.addmatmat_host(a[20:500], 480, k, c0[20:500].copy(), n, 1.3, 2.0)
    assert dev.last_product()["kernel"].split("<")[0] == family
    assert_bits_equal(part, got[20:500], "m = 480 slice vs rows 20..499 of the m = 512 product (same kernel family)")
    pa = torch.from_numpy(a).pin_memory(); pc = torch.from_numpy(c0.copy()).pin_memory()
    dev.addmatmat_host(pa, m, k, pc, n, 1.3, 2.0)
    assert _sampled_backward_error(pc.numpy()[:, rows], codes[rows], tab, a, c0[:, rows], 1.3, 2.0) <= TOL
    tile = _with_env("SBLAS_B200_SPMM_TC", "0", lambda: dev.addmatmat_host(a, m, k, c0.copy(), n, 1.3, 2.0))
    assert dev.last_product()["kernel"].startswith("spmm_tile")
    _with_env("SBLAS_B200_SPMM_TC", "0", lambda: dev.addmatmat_host(pa, m, k, pc.copy_(torch.from_numpy(c0)), n, 1.3, 2.0))
    assert_bits_equal(pc.numpy(), tile, "pinned + sliced host route at full size (tile kernels)")
    # row slabs (8 virtual shards) concatenate to the same bits
    pos, code, table, T, bb, bib = S.members()
    splits = sb.partition(bb, bib, n, 8)
    for g in (0, 3, 7):
        lo, hi = int(splits[g]), int(splits[g + 1])
        slab = S.upload_slab(lo, hi)
        cs = np.ascontiguousarray(c0[:, lo:hi])
        ys = slab.addmatmat_host(a, m, k, cs, hi - lo, 1.3, 2.0)
        assert_bits_equal(ys, got[:, lo:hi], f"slab {g}")


def test_config4_full_size_sampled_rows():
    # BASELINE.json configs[3]: power-law SpMV, 2^23 rows, ~16 nnz/row, direct CSR
    n = k = 1 << 23
    row_ptr, col, code = gen.powerlaw_csr(n, k, seed=14)
    tab = gen.codebook()
    dev = sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k)
    assert dev.nnz == int(row_ptr[-1])
    a = gen.dense_uniform((1, k), 2); c0 = gen.dense_uniform((1, n), 3)
    got = dev.addmatmat_host(a, 1, k, c0.copy(), n, 1.3, 2.0).reshape(-1)
    lens = np.diff(row_ptr)
    rng = np.random.default_rng(6)
    rows = np.concatenate([rng.choice(n, 2000, replace=False), np.argsort(lens)[-20:], np.flatnonzero(lens == lens.min())[:20]])
    tab64 = tab.astype(np.float64); a64 = a.reshape(-1).astype(np.float64)
    worst = 0.0
    for j in rows:
        s, e = int(row_ptr[j]), int(row_ptr[j + 1])
        w = tab64[code[s:e]]; xv = a64[col[s:e]]
        y64 = 2.0 * float(c0[0, j]) + 1.3 * float(np.dot(w, xv))
        bound = abs(2.0 * float(c0[0, j])) + 1.3 * float(np.dot(np.abs(w), np.abs(xv)))
        worst = max(worst, abs(float(got[j]) - y64) / max(bound, 1e-300))
    assert worst <= TOL, worst
    again = dev.addmatmat_host(a, 1, k, c0.copy(), n, 1.3, 2.0).reshape(-1)
    assert_bits_equal(again, got, "run-to-run determinism at full size")


def test_config5_full_size_dense_spmv_sampled_rows_and_row_kernel():
    # BASELINE.json configs[4], large end of the sweep: 16384 x 16384 at 50 % sparsity, m = 1 (dense-code kernel)
    n = k = 16384
    codes = gen.codes_bernoulli(n, k, 0.50, seed=100)
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, 1)
    dev = S.device_matrix()
    assert dev.nnz == int((codes != 255).sum())
    a = gen.dense_uniform((1, k), 2); c0 = gen.dense_uniform((1, n), 3)
    before = dev.info().device_bytes
    got = dev.addmatmat_host(a, 1, k, c0.copy(), n, 1.3, 2.0)
    assert dev.info().device_bytes >= before + n * k           # the dense code matrix was built: that kernel ran
    rows = np.random.default_rng(7).choice(n, 64, replace=False)
    err = _sampled_backward_error(got[:, rows], codes[rows], tab, a, c0[:, rows], 1.3, 2.0)
    assert err <= TOL, err
    again = dev.addmatmat_host(a, 1, k, c0.copy(), n, 1.3, 2.0)
    assert_bits_equal(again, got, "run-to-run determinism at full size")
    # the CSR row kernels on the same handle agree to rounding (another summation order, same contract)
    rows_k = _with_env("SBLAS_B200_SPMV_DENSE", "0", lambda: dev.addmatmat_host(a, 1, k, c0.copy(), n, 1.3, 2.0))
    assert _sampled_backward_error(rows_k[:, rows], codes[rows], tab, a, c0[:, rows], 1.3, 2.0) <= TOL
    assert np.max(np.abs(rows_k - got)) <= 1e-5 * float(np.max(np.abs(got))) * 16
    # linearity in x: W (x1 + x2) = W x1 + W x2 up to rounding (beta = 0 is not "overwrite": C = 0 going in)
    a2 = gen.dense_uniform((1, k), 9)
    z = np.zeros((1, n), np.float32)
    y1 = dev.addmatmat_host(a, 1, k, z.copy(), n, 1.0, 1.0); y2 = dev.addmatmat_host(a2, 1, k, z.copy(), n, 1.0, 1.0)
    y12 = dev.addmatmat_host(a + a2, 1, k, z.copy(), n, 1.0, 1.0)
    scale = float(np.abs(a).sum() + np.abs(a2).sum()) / k * 64.0          # ~ sum |w||x| of one row
    assert np.max(np.abs(y12 - (y1 + y2))) <= 1e-5 * scale * 64


def test_products_of_one_handle_on_two_streams_are_ordered_by_the_library():
    """The Xt staging buffer belongs to the handle; products enqueued on different streams must be
    ordered by the library (event), not race on it."""
    import torch
    n, k, m = 2048, 2048, 96
    codes = gen.codes_bernoulli(n, k, 0.05, seed=81)
    tab = gen.codebook()
    dev = upload(O.copyform(codes, n, k, k, tab, 255, True))
    a1 = torch.from_numpy(gen.dense_uniform((m, k), 1)).cuda(); a2 = torch.from_numpy(gen.dense_uniform((m, k), 2)).cuda()
    want1 = torch.zeros((m, n), device="cuda"); want2 = torch.zeros((m, n), device="cuda")
    dev.addmatmat_dev(a1, m, k, want1, n, 1.0, 0.0); dev.addmatmat_dev(a2, m, k, want2, n, 1.0, 0.0)
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(20):
        c1 = torch.zeros((m, n), device="cuda"); c2 = torch.zeros((m, n), device="cuda")
        torch.cuda.synchronize()
        dev.addmatmat_dev(a1, m, k, c1, n, 1.0, 0.0, stream=s1.cuda_stream)
        dev.addmatmat_dev(a2, m, k, c2, n, 1.0, 0.0, stream=s2.cuda_stream)
        torch.cuda.synchronize()
        assert torch.equal(c1, want1) and torch.equal(c2, want2)


def test_narrow_tiles_for_small_m_on_a_large_matrix():
    """m <= 16 takes the 16-column tile (two entries of a slot per step), m <= 8 on a matrix of >= 6 Mi
    non-zeros the 8-column tile (four per step): sampled rows against fp64, and one output element
    does not depend on which other A rows are in the call within the same tile width."""
    n = k = 8192
    codes = gen.codes_bernoulli(n, k, 0.10, seed=91)           # 6.7 M non-zeros
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, 1)
    dev = S.device_matrix()
    assert dev.nnz >= (6 << 20)
    rows = np.random.default_rng(7).choice(n, 40, replace=False)
    for m in (2, 5, 8, 9, 16):
        a = gen.dense_uniform((m, k), 90 + m); c0 = gen.dense_uniform((m, n + 4), 91 + m)
        got = dev.addmatmat_host(a, m, k, c0.copy(), n + 4, 1.3, 2.0)
        err = _sampled_backward_error(got[:, rows], codes[rows], tab, a, c0[:, rows], 1.3, 2.0)
        assert err <= TOL, (m, err)
        again = dev.addmatmat_host(a, m, k, c0.copy(), n + 4, 1.3, 2.0)
        assert_bits_equal(again, got, f"determinism m={m}")
    a = gen.dense_uniform((8, k), 3); c0 = gen.dense_uniform((8, n), 4)
    full = dev.addmatmat_host(a, 8, k, c0.copy(), n, 1.0, 1.0)
    part = dev.addmatmat_host(a[2:5], 3, k, c0[2:5].copy(), n, 1.0, 1.0)
    assert_bits_equal(part, full[2:5], "m = 3 vs rows 2..4 of m = 8 (same 8-column tile)")


# ------------------------------------------------------------------ batched small-m calls in a CUDA graph
def test_products_can_be_captured_into_a_cuda_graph_and_replayed():
    # A caller with many small products (m = 1 .. 8: launch bound, a few us each) records them once into a
    # CUDA graph and replays the whole batch with one launch.  The device entry point must therefore be
    # capturable after one warm-up call per shape (workspaces allocated, kernels configured): no allocation,
    # no synchronisation, no host-side state that a replay would miss.  Results: bit-identical to eager calls.
    import torch
    n, k = 3000, 1100
    codes = gen.codes_bernoulli(n, k, 0.05, seed=91)
    S = sb.SparseMatrix(codes, n, k, k, gen.codebook(), 255, 1)
    dev = S.device_matrix()
    cuda = torch.device("cuda", 0)
    ms = (1, 3, 8, 40)
    a = {m: torch.from_numpy(gen.dense_uniform((m, k), 100 + m)).to(cuda) for m in ms}
    c0 = {m: torch.from_numpy(gen.dense_uniform((m, n), 200 + m)).to(cuda) for m in ms}
    want = {}
    for m in ms:                                        # eager: also the warm-up of every shape
        c = c0[m].clone()
        dev.addmatmat_dev(a[m], m, k, c, n, 1.3, 0.7)
        dev.addmatmat_dev(a[m], m, k, c, n, 1.0, 1.0)
        torch.cuda.synchronize()
        want[m] = c.cpu().numpy()
    c = {m: c0[m].clone() for m in ms}
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    launches0 = sb.lib().sblas_b200_launch_count()
    with torch.cuda.graph(graph, stream=side):
        for m in ms:
            dev.addmatmat_dev(a[m], m, k, c[m], n, 1.3, 0.7, stream=side.cuda_stream)
            dev.addmatmat_dev(a[m], m, k, c[m], n, 1.0, 1.0, stream=side.cuda_stream)
    assert sb.lib().sblas_b200_launch_count() > launches0
    for rep in range(2):
        for m in ms:
            c[m].copy_(c0[m])
        graph.replay()
        torch.cuda.synchronize()
        for m in ms:
            assert_bits_equal(c[m].cpu().numpy(), want[m], f"graph replay {rep} m={m}")
    # and the handle still works eagerly afterwards, on another stream
    c1 = c0[8].clone()
    dev.addmatmat_dev(a[8], 8, k, c1, n, 1.3, 0.7)
    dev.addmatmat_dev(a[8], 8, k, c1, n, 1.0, 1.0)
    torch.cuda.synchronize()
    assert_bits_equal(c1.cpu().numpy(), want[8], "eager after capture")


# ------------------------------------------------------------------ column-blocked two-pass SpMV (spmv_blocked.cu)
def _blocked_upload(row_ptr, col, code, tab, n, k, cbw, rs, long_threshold=None):
    def up():
        return sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k)
    fn = lambda: _with_env("SBLAS_B200_BLOCKED_CBW", str(cbw), lambda: _with_env("SBLAS_B200_BLOCKED_RS", str(rs), up))
    if long_threshold is not None:
        inner = fn
        fn = lambda: _with_env("SBLAS_B200_LONG_THRESHOLD", str(long_threshold), inner)
    return _with_env("SBLAS_B200_SPMV_BLOCKED", "1", fn)


@pytest.mark.parametrize("n,k,cbw,rs,x_min,cap,long_threshold", [
    (5000, 3001, 256, 37, 4, 900, None),        # tail column block, runs of equal rows across lanes, n not a multiple of rs
    (70001, 40000, 4096, 100, 8, 30000, 500),   # long rows (split-row path) beside the blocked ones
    (300, 100000, 32768, 1, 2, 64, None),       # one row per sub-block, default tile width
    (4097, 513, 64, 1792, 16, 513, None),       # widest sub-block, rows as long as k
])
def test_blocked_spmv_against_the_oracle_and_the_row_kernels(n, k, cbw, rs, x_min, cap, long_threshold):
    import torch
    row_ptr, col, code = gen.powerlaw_csr(n, k, seed=77, x_min=x_min, cap=cap)
    lens = np.diff(row_ptr).astype(np.int64)                   # a few rows lose their entries, incl. the last one
    keep = np.ones(col.shape[0], bool)
    for j in (0, 5, 6, n - 1):
        keep[row_ptr[j]:row_ptr[j + 1]] = False; lens[j] = 0
    col, code = col[keep], code[keep]
    row_ptr = np.concatenate(([0], np.cumsum(lens))).astype(np.int32)
    tab = gen.codebook()
    dev = _blocked_upload(row_ptr, col, code, tab, n, k, cbw, rs, long_threshold=long_threshold)
    plain = _with_env("SBLAS_B200_SPMV_BLOCKED", "0", lambda: sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k))
    if long_threshold:
        assert dev.info().long_rows > 0
    a = gen.dense_uniform((1, k), 1); c0 = gen.dense_uniform((1, n), 2); c0[0, 5] = -0.0
    for alpha, beta in ((1.0, 1.0), (1.3, 2.0), (0.7, 0.0)):
        got = dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta)
        lp = dev.last_product()
        assert lp["kernel"].startswith("spmv_blocked") and lp["m"] == 1 and lp["algorithmic_flops"] == 2 * int(row_ptr[-1]), lp
        y64, bound = O.addmatmat_csr_f64(row_ptr, col, code, tab, 255, n, k, a, 1, k, c0, n, alpha, beta)
        back = O.error_measures(got, 1, n, n, y64, bound)[0]
        want = O.addmatmat_csr(row_ptr, col, code, tab, 255, n, k, a, 1, k, c0.copy(), n, alpha, beta)
        rback = O.error_measures(want, 1, n, n, y64, bound)[0]
        assert back <= TOL, (alpha, beta, back, rback)
        # rows without entries: exactly the scaled value, sign of zero kept (the reference's own bits)
        for j in (0, 5, 6, n - 1):
            assert got.reshape(-1)[j].tobytes() == want.reshape(-1)[j].tobytes(), (j, alpha, beta)
        other = plain.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta)
        assert O.error_measures(other, 1, n, n, y64, bound)[0] <= TOL
        assert not plain.last_product()["kernel"].startswith("spmv_blocked")
        assert_bits_equal(dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta), got, "run-to-run determinism")
    # x off the 16-byte grid (scalar tile loads) and y declared peer memory (sums added by reductions, y never read)
    L = sb.lib()
    xbuf = torch.zeros(k + 1, device="cuda"); xbuf[1:] = torch.from_numpy(a.reshape(-1)).cuda()
    for alpha, beta in ((1.0, 1.0), (1.3, 2.0)):
        want = torch.from_numpy(c0.copy()).cuda()
        dev.addmatmat_dev(torch.from_numpy(a).cuda(), 1, k, want, n, alpha, beta)
        got = torch.from_numpy(c0.copy()).cuda()
        dev.addmatmat_dev(xbuf[1:], 1, k, got, n, alpha, beta)
        torch.cuda.synchronize()
        assert_bits_equal(got.cpu().numpy(), want.cpu().numpy(), "misaligned x")
        buf = torch.from_numpy(c0.copy()).cuda()
        assert L.sblas_b200_peer_register(C.c_void_p(buf.data_ptr()), buf.numel() * 4) == _capi.OK
        try:
            dev.addmatmat_dev(torch.from_numpy(a).cuda(), 1, k, buf, n, alpha, beta)
            torch.cuda.synchronize()
        finally:
            assert L.sblas_b200_peer_unregister(C.c_void_p(buf.data_ptr())) == _capi.OK
        assert_bits_equal(buf.cpu().numpy(), want.cpu().numpy(), f"peer range alpha={alpha} beta={beta}")


def test_blocked_spmv_is_built_only_on_request_for_small_matrices():
    # a band: every row's entries share one column block -> the row kernels' gathers hit L1, the second copy would lose
    n = k = 1 << 16
    lens = np.full(n, 8, np.int64)
    row_ptr = np.concatenate(([0], np.cumsum(lens))).astype(np.int32)
    col = (np.repeat(np.arange(n, dtype=np.int64), 8) + np.tile(np.arange(8, dtype=np.int64), n)) % k
    col = np.sort(col.reshape(n, 8), axis=1).reshape(-1).astype(np.int32)
    code = (np.arange(col.shape[0]) % 255).astype(np.uint8)
    tab = gen.codebook()
    dev = _with_env("SBLAS_B200_SPMV_BLOCKED", "1", lambda: sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k))
    a = gen.dense_uniform((1, k), 1)
    dev.addmatmat_host(a, 1, k, np.zeros((1, n), np.float32), n, 1.0, 1.0)
    assert dev.last_product()["kernel"].startswith("spmv_blocked")          # forced: built whatever the shape
    dev = sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k)
    dev.addmatmat_host(a, 1, k, np.zeros((1, n), np.float32), n, 1.0, 1.0)
    assert dev.last_product()["kernel"].startswith("spmv_rows")              # small: the row kernels stay


def test_host_entry_point_cuts_a_long_y_into_column_ranges():
    """m = 1 with y of 8 MB or more and caller buffers from sblas_b200_host_alloc: x in, the column ranges of y in, ONE product,
    then range after range combined and copied out (product.cu: host_product_overlapped, colmode).  Same kernels as from
    pageable buffers, so the same bits; columns of S without entries keep their (scaled) value bit for bit."""
    n, k = (5 << 19) + 37, 4096                                 # 10 MB of y, not a multiple of the range alignment
    row_ptr, col, code = gen.powerlaw_csr(n, k, seed=23, x_min=2, cap=64)
    lens = np.diff(row_ptr)
    tab = gen.codebook()
    dev = sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k)
    a = gen.dense_uniform((1, k), 2); c0 = gen.dense_uniform((1, n), 3)
    empty = np.flatnonzero(lens == 0)[:8]
    if empty.size: c0[0, empty[0]] = -0.0
    for alpha, beta in ((1.0, 1.0), (1.3, 2.0), (0.7, 0.0)):
        want = dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta)              # numpy (pageable) buffers
        h_a = sb.host_buffer((1, k)); h_a[...] = a
        h_c = sb.host_buffer((1, n)); h_c[...] = c0
        got = dev.addmatmat_host(h_a, 1, k, h_c, n, alpha, beta)
        assert_bits_equal(np.asarray(got).reshape(-1), want.reshape(-1), f"column ranges vs one pass, alpha={alpha} beta={beta}")
    dev.free()


def test_spmv_with_long_rows_can_be_captured_into_a_cuda_graph():
    # the long rows' segment sums run on a side stream of the handle (forked from and joined back into the caller's stream):
    # inside a capture that fork and join become graph edges; replays give the eager bits
    import torch
    n, k = 70000, 30000
    row_ptr, col, code = gen.powerlaw_csr(n, k, seed=31, x_min=4, cap=20000)
    dev = sb.DeviceMatrix.upload_csr(row_ptr, col, code, gen.codebook(), 255, n, k)
    assert dev.info().long_rows > 0, "the case needs rows on the split-row path"
    cuda = torch.device("cuda", 0)
    x = torch.from_numpy(gen.dense_uniform((1, k), 5)).to(cuda)
    y0 = torch.from_numpy(gen.dense_uniform((1, n), 6)).to(cuda)
    y = y0.clone()
    dev.addmatmat_dev(x, 1, k, y, n, 1.3, 0.7)                  # eager: warm-up (side stream, workspaces) and the expected bits
    torch.cuda.synchronize()
    want = y.cpu().numpy()
    yg = y0.clone()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        dev.addmatmat_dev(x, 1, k, yg, n, 1.3, 0.7, stream=side.cuda_stream)
    for rep in range(2):
        yg.copy_(y0)
        graph.replay()
        torch.cuda.synchronize()
        assert_bits_equal(yg.cpu().numpy(), want, f"replay {rep}")
    dev.free()
