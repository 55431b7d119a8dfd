"""K-spmm-tc (src/sparse/spmm_tc.cu): the tensor-core form of the product (tcgen05, TF32 x 3 split, fp32 accumulators in
tensor memory) against the oracle.  Same contract as every other kernel: backward error <= 1e-5 against fp64, rows of W
without entries only get the scale, padding columns of C untouched, run-to-run identical; non-finite x must not leak
through absent entries (the reference never touches x there)."""
import os

import numpy as np
import pytest

import sparsematrix_b200 as sb
from sparsematrix_b200 import gen
from oracle import oracle as O
from helpers import assert_bits_equal

pytestmark = pytest.mark.gpu

TOL = 1e-5


def _env(name, value, fn):
    old = os.environ.get(name)
    os.environ[name] = value
    try:
        return fn()
    finally:
        if old is None: os.environ.pop(name, None)
        else: os.environ[name] = old


def _check(dev, s, m, alpha, beta, pad_a=0, pad_c=0, seed=0, expect_tc=True):
    n, k = s.cols, s.rows
    lda, ldc = k + pad_a, n + pad_c
    a = gen.dense_uniform((m, lda), 10 + seed)
    c0 = gen.dense_uniform((m, ldc), 20 + seed)
    got = _env("SBLAS_B200_SPMM_TC", "1", lambda: dev.addmatmat_host(a, m, lda, c0.copy(), ldc, alpha, beta))
    kernel = dev.last_product()["kernel"]
    assert kernel.startswith("spmm_tc") == expect_tc, kernel
    y64, bound = O.addmatmat_f64(s, a, m, lda, c0, ldc, alpha, beta)
    back, norm, _ = O.error_measures(got, m, n, ldc, y64, bound)
    want = O.addmatmat(s, a, m, lda, c0.copy(), ldc, alpha, beta)
    rback, _, _ = O.error_measures(want, m, n, ldc, y64, bound)
    assert back <= TOL, f"{kernel} m={m} alpha={alpha} beta={beta}: backward error {back:.3e} (reference arithmetic {rback:.3e})"
    if pad_c:
        assert_bits_equal(got.reshape(m, ldc)[:, n:], c0.reshape(m, ldc)[:, n:], "ldc padding")
    again = _env("SBLAS_B200_SPMM_TC", "1", lambda: dev.addmatmat_host(a, m, lda, c0.copy(), ldc, alpha, beta))
    assert_bits_equal(again, got, "run-to-run determinism")
    return got, want, back


@pytest.mark.parametrize("n,k,dens", [(300, 257, 0.3), (1024, 1024, 0.1), (700, 1300, 0.5)])
def test_tensor_core_spmm_against_the_oracle(n, k, dens):
    codes = gen.codes_bernoulli(n, k, dens, seed=n + k)
    codes[5, :] = 255; codes[n - 1, :] = 255                   # rows of W without entries
    tab = gen.codebook(255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    worst = 0.0
    for m, alpha, beta, pa, pc in ((32, 1.0, 1.0, 0, 0), (37, 1.3, 2.0, 3, 5), (64, 1.0, 0.0, 0, 0), (100, 1.8, 1.2, 0, 1),
                                   (128, 1.0, 1.0, 0, 0), (200, 0.7, 1.0, 1, 0), (300, 1.0, 1.0, 0, 0)):
        got, want, back = _check(dev, s, m, alpha, beta, pa, pc, seed=m)
        worst = max(worst, back)
        for j in (5, n - 1):                                   # no entry: exactly the reference's bits (scaled, sign of zero kept)
            assert_bits_equal(got.reshape(m, -1)[:, j], want.reshape(m, -1)[:, j], f"empty row {j}")
    print(f"spmm_tc {n}x{k} at {dens}: worst backward error {worst:.2e}")
    dev.free()


def test_tensor_core_spmm_split_k_is_deterministic_and_equal_in_tolerance():
    n, k, m = 256, 4096, 96                                    # 2 row blocks x 1 m tile: the k range is split over CTAs
    codes = gen.codes_bernoulli(n, k, 0.2, seed=61)
    s = O.copyform(codes, n, k, k, gen.codebook(255, 1), 255, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    got, _, _ = _check(dev, s, m, 1.3, 2.0, seed=1)
    assert int(dev.last_product()["kernel"].split(">x")[1]) > 1, dev.last_product()["kernel"]       # really split
    for ks in ("1", "3", "16"):
        other, _, _ = _env("SBLAS_B200_TC_KSPLIT", ks, lambda: _check(dev, s, m, 1.3, 2.0, seed=1))
        y64, bound = O.addmatmat_f64(s, gen.dense_uniform((m, k), 11), m, k, gen.dense_uniform((m, n), 21), n, 1.3, 2.0)
        assert float(np.max(np.abs(other.astype(np.float64) - got.astype(np.float64)) / np.maximum(bound, 1e-300))) <= TOL
    dev.free()


def test_tensor_core_spmm_shares_added_by_their_own_kernel_or_by_the_last_cta_give_the_same_bits():
    # grids of one wave add the shares of k with tc_fixup_kernel, larger ones inside the contraction (the last CTA of a
    # tile); SBLAS_B200_TC_FIXUP=0 forces the latter.  Same sums in the same order.
    n, k = 700, 2048                                           # 6 row blocks x m tiles x 4 shares: one wave up to m = 256 x 4
    codes = gen.codes_bernoulli(n, k, 0.3, seed=77)
    codes[9, :] = 255
    s = O.copyform(codes, n, k, k, gen.codebook(255, 1), 255, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    for m, alpha, beta, pc in ((40, 1.0, 1.0, 0), (128, 1.7, 0.0, 3), (300, 0.9, 2.0, 1)):
        own, _, _ = _check(dev, s, m, alpha, beta, 0, pc, seed=m)
        assert int(dev.last_product()["kernel"].split(">x")[1]) > 1
        inside, _, _ = _env("SBLAS_B200_TC_FIXUP", "0", lambda: _check(dev, s, m, alpha, beta, 0, pc, seed=m))
        assert_bits_equal(own, inside, f"m={m}: shares added by tc_fixup_kernel vs by the last CTA")
    # more shares than the last CTA's staging area holds at a time (4 or 6 slots), shares that are not multiples of it
    for ks in ("7", "13"):
        def both():
            own, _, _ = _check(dev, s, 128, 1.3, 1.0, 0, 0, seed=3)
            inside, _, _ = _env("SBLAS_B200_TC_FIXUP", "0", lambda: _check(dev, s, 128, 1.3, 1.0, 0, 0, seed=3))
            return own, inside
        own, inside = _env("SBLAS_B200_TC_KSPLIT", ks, both)
        assert dev.last_product()["kernel"].endswith("x" + ks), dev.last_product()["kernel"]
        assert_bits_equal(own, inside, f"{ks} shares: tc_fixup_kernel vs the last CTA")
    # non-finite x with the separate kernel: the fallback computes the product, nothing of the tensor-core pass is stored
    a = gen.dense_uniform((64, k), 5)
    a[3, 17] = np.inf
    c0 = gen.dense_uniform((64, n), 6)
    got = _env("SBLAS_B200_SPMM_TC", "1", lambda: dev.addmatmat_host(a, 64, k, c0.copy(), n, 1.0, 1.0))
    want = _env("SBLAS_B200_SPMM_TC", "0", lambda: dev.addmatmat_host(a, 64, k, c0.copy(), n, 1.0, 1.0))
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin)
    np.testing.assert_allclose(got[fin], want[fin], rtol=1e-4, atol=1e-4)
    dev.free()


def test_tensor_core_spmm_never_lets_x_through_where_no_entry_is():
    n, k, m = 256, 512, 40
    codes = gen.codes_bernoulli(n, k, 0.25, seed=71)
    codes[:, 7] = 255                                          # no W row uses x column 7 ...
    codes[3, 9] = 17                                           # ... and row 3 uses column 9
    tab = gen.codebook(255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    a = gen.dense_uniform((m, k), 1); c0 = gen.dense_uniform((m, n), 2)
    a[2, 7] = np.inf; a[4, 7] = np.nan                         # must not reach any output
    got = _env("SBLAS_B200_SPMM_TC", "1", lambda: dev.addmatmat_host(a, m, k, c0.copy(), n, 1.0, 1.0))
    assert dev.last_product()["kernel"].startswith("spmm_tc")
    assert np.isfinite(got).all()
    a_fin = a.copy(); a_fin[2, 7] = 0.0; a_fin[4, 7] = 0.0
    y64, bound = O.addmatmat_f64(s, a_fin, m, k, c0, n, 1.0, 1.0)
    assert O.error_measures(got, m, n, n, y64, bound)[0] <= TOL
    a[6, 9] = np.inf                                           # this one is used by row 3: C[6][3] becomes inf, like the reference's
    got = _env("SBLAS_B200_SPMM_TC", "1", lambda: dev.addmatmat_host(a, m, k, c0.copy(), n, 1.0, 1.0))
    want = O.addmatmat(s, a, m, k, c0.copy(), n, 1.0, 1.0)
    assert np.array_equal(np.isfinite(got), np.isfinite(want))
    # a finite call afterwards takes the tensor-core path again and is right
    got = _env("SBLAS_B200_SPMM_TC", "1", lambda: dev.addmatmat_host(a_fin, m, k, c0.copy(), n, 1.0, 1.0))
    assert O.error_measures(got, m, n, n, y64, bound)[0] <= TOL
    dev.free()


def test_tensor_core_path_is_chosen_by_density_and_m():
    tab = gen.codebook(255, 1)
    for dens, m, expect in ((0.25, 128, True), (0.25, 16, False), (0.02, 128, False)):
        n = k = 4096
        codes = gen.codes_bernoulli(n, k, dens, seed=81)
        S = sb.SparseMatrix(codes, n, k, k, tab, 255, sb.SparseMatrix.Trans)
        dev = S.device_matrix()
        a = gen.dense_uniform((m, k), 1); c0 = gen.dense_uniform((m, n), 2)
        dev.addmatmat_host(a, m, k, c0.copy(), n, 1.0, 1.0)
        assert dev.last_product()["kernel"].startswith("spmm_tc") == expect, (dens, m, dev.last_product()["kernel"])


def test_tensor_core_spmm_into_a_registered_peer_range_adds_with_reductions():
    """C declared peer memory (sblas_b200_peer_register) and beta == 1: the epilogue adds at the owner (red.global.add) and never
    reads C; same bits as the local read-modify-write (one rounding either way), rows without entries untouched."""
    import ctypes as C
    import torch
    from sparsematrix_b200 import _capi
    n, k, m = 384, 640, 96
    codes = gen.codes_bernoulli(n, k, 0.3, seed=91)
    codes[7, :] = 255
    s = O.copyform(codes, n, k, k, gen.codebook(255, 1), 255, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    a = torch.from_numpy(gen.dense_uniform((m, k), 1)).cuda()
    c0 = gen.dense_uniform((m, n), 2); c0[3, 7] = -0.0
    L = sb.lib()
    # the shares of k added by tc_fixup_kernel ("1": the grid is one wave) and by the last CTA of a tile ("0"): both epilogues
    for fixup in ("1", "0"):
        for beta in (1.0, 0.5):
            want = torch.from_numpy(c0.copy()).cuda()
            _env("SBLAS_B200_TC_FIXUP", fixup, lambda: _env("SBLAS_B200_SPMM_TC", "1", lambda: dev.addmatmat_dev(a, m, k, want, n, 1.3, beta)))
            assert dev.last_product()["kernel"].startswith("spmm_tc")
            assert int(dev.last_product()["kernel"].split(">x")[1]) > 1
            buf = torch.from_numpy(c0.copy()).cuda()
            assert L.sblas_b200_peer_register(C.c_void_p(buf.data_ptr()), buf.numel() * 4) == _capi.OK
            try:
                _env("SBLAS_B200_TC_FIXUP", fixup, lambda: _env("SBLAS_B200_SPMM_TC", "1", lambda: dev.addmatmat_dev(a, m, k, buf, n, 1.3, beta)))
                torch.cuda.synchronize()
            finally:
                assert L.sblas_b200_peer_unregister(C.c_void_p(buf.data_ptr())) == _capi.OK
            assert dev.last_product()["kernel"].startswith("spmm_tc")
            assert_bits_equal(buf.cpu().numpy(), want.cpu().numpy(), f"peer range beta={beta} fixup={fixup}")
    dev.free()


def test_tensor_core_spmm_products_of_different_shapes_share_one_workspace():
    """Many tiles (n = 25 600: 200 row blocks) and m going up and down on one handle: the tile counters of the wide product
    must not lie where the narrow one kept its X images (a layout bug of the first version, caught by bench.py's check)."""
    n, k = 25600, 256
    codes = gen.codes_bernoulli(n, k, 0.3, seed=95)
    s = O.copyform(codes, n, k, k, gen.codebook(255, 1), 255, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    for m in (256, 512, 256, 64, 512):
        _check(dev, s, m, 1.0, 1.0, seed=m)
    dev.free()


def test_tensor_core_spmm_can_be_captured_into_a_cuda_graph():
    """Like the other kernels (tests/test_gpu_parity.py, graphs): after one eager call per shape the product records into a
    CUDA graph and replays with the same bits; the call id of the non-finite-x flag is baked into the recording, which is fine."""
    import torch
    n, k, m = 512, 1024, 128
    codes = gen.codes_bernoulli(n, k, 0.3, seed=97)
    s = O.copyform(codes, n, k, k, gen.codebook(255, 1), 255, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    a = torch.from_numpy(gen.dense_uniform((m, k), 1)).cuda()
    c0 = torch.from_numpy(gen.dense_uniform((m, n), 2)).cuda()
    os.environ["SBLAS_B200_SPMM_TC"] = "1"
    try:
        want = c0.clone()
        dev.addmatmat_dev(a, m, k, want, n, 1.3, 0.7)
        torch.cuda.synchronize()
        assert dev.last_product()["kernel"].startswith("spmm_tc")
        c = c0.clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            dev.addmatmat_dev(a, m, k, c, n, 1.3, 0.7, stream=side.cuda_stream)
        assert dev.last_product()["kernel"].startswith("spmm_tc")
        for rep in range(3):
            c.copy_(c0)
            graph.replay()
            torch.cuda.synchronize()
            assert_bits_equal(c.cpu().numpy(), want.cpu().numpy(), f"graph replay {rep}")
    finally:
        os.environ.pop("SBLAS_B200_SPMM_TC", None)
    dev.free()
