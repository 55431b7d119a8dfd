"""K-spmv (native-stream form, src/sparse/spmv_native.cu): y = beta*y + alpha * W x multiplied straight from the reference's
pos_index_ / val_index_ streams (the walk of /root/reference/src/sparse/kernel.cc:780-796, the filler rule of
sparse-matrix.cc:46-51).  Same contract as every other kernel: backward error <= 1e-5 against fp64, columns of S without
entries only get the scale (bit for bit the reference's value), run-to-run identical, x never read where no entry is."""
import os

import numpy as np
import pytest

import sparsematrix_b200 as sb
from sparsematrix_b200 import gen
from oracle import oracle as O
from helpers import assert_bits_equal

pytestmark = pytest.mark.gpu

TOL = 1e-5


class _Env:
    def __init__(self, **kv):
        self.kv, self.old = kv, {}

    def __enter__(self):
        for k, v in self.kv.items():
            self.old[k] = os.environ.get(k)
            os.environ[k] = v

    def __exit__(self, *exc):
        for k, v in self.old.items():
            if v is None: os.environ.pop(k, None)
            else: os.environ[k] = v


def _upload_native(s, **kw):
    with _Env(SBLAS_B200_KEEP_NATIVE="1"):
        return sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols, **kw)


def _product(dev, a, k, c0, n, alpha, beta, native=True):
    with _Env(SBLAS_B200_SPMV_NATIVE="1" if native else "0", SBLAS_B200_SPMV_DENSE="0"):
        got = dev.addmatmat_host(a, 1, k, c0.copy(), n, alpha, beta)
    assert dev.last_product()["kernel"].startswith("spmv_native") == native, dev.last_product()["kernel"]
    return got


@pytest.mark.parametrize("n,k,dens", [(700, 1300, 0.10), (1024, 4096, 0.02), (300, 9000, 0.003), (513, 777, 0.6), (256, 256, 1.0)])
def test_native_stream_spmv_against_the_oracle(n, k, dens):
    codes = gen.codes_bernoulli(n, k, dens, seed=n + k)
    codes[5, :] = 255; codes[n - 1, :] = 255                   # columns of S (rows of W) without entries
    if n > 600: codes[256:512, :] = 255                         # a whole panel without entries: not recorded at all
    tab = gen.codebook(255, 1)
    s = O.copyform(codes, n, k, k, tab, 255, True)
    assert dens > 0.01 or int((s.code >= s.T).sum()) > 0        # the sparse cases carry fillers
    dev = _upload_native(s)
    for alpha, beta in ((1.0, 1.0), (1.3, 2.0), (0.7, 0.0)):
        a = gen.dense_uniform((1, k), 10); c0 = gen.dense_uniform((1, n), 20); c0[0, 5] = -0.0
        got = _product(dev, a, k, c0, n, alpha, beta)
        y64, bound = O.addmatmat_f64(s, a, 1, k, c0, n, alpha, beta)
        back, _, _ = O.error_measures(got, 1, n, n, y64, bound)
        assert back <= TOL, (alpha, beta, back)
        want = O.addmatmat(s, a, 1, k, c0.copy(), n, alpha, beta)
        for j in (5, n - 1) + ((300,) if n > 600 else ()):      # no entry: exactly the reference's bits (scaled, -0 kept)
            assert_bits_equal(got.reshape(-1)[j:j + 1], want.reshape(-1)[j:j + 1], f"empty column {j}")
        assert_bits_equal(_product(dev, a, k, c0, n, alpha, beta), got, "run-to-run determinism")
        other = _product(dev, a, k, c0, n, alpha, beta, native=False)
        assert float(np.max(np.abs(other.astype(np.float64) - got.astype(np.float64)) / np.maximum(bound, 1e-300))) <= TOL
    dev.free()


def test_native_stream_spmv_never_reads_x_where_no_entry_is():
    n, k = 600, 2000
    codes = gen.codes_bernoulli(n, k, 0.05, seed=3)
    codes[:, 77] = 255                                          # no W row uses x[77] ...
    codes[10, 77] = 3                                           # ... except row 10
    s = O.copyform(codes, n, k, k, gen.codebook(255, 1), 255, True)
    dev = _upload_native(s)
    a = gen.dense_uniform((1, k), 1); a[0, 77] = np.inf
    c0 = gen.dense_uniform((1, n), 2)
    got = _product(dev, a, k, c0, n, 1.0, 1.0).reshape(-1)
    assert not np.isfinite(got[10]) and np.isfinite(np.delete(got, 10)).all()
    dev.free()


def test_native_stream_spmv_of_a_slab_and_into_a_registered_peer_range():
    import ctypes as C
    import torch
    from sparsematrix_b200 import _capi
    n, k = 1100, 3000
    codes = gen.codes_bernoulli(n, k, 0.08, seed=8)
    codes[700, :] = 255
    s = O.copyform(codes, n, k, k, gen.codebook(255, 1), 255, True)
    a = gen.dense_uniform((1, k), 4); c0 = gen.dense_uniform((1, n), 5)
    whole = _upload_native(s)
    full = _product(whole, a, k, c0, n, 1.3, 2.0).reshape(-1)
    for lo, hi in ((0, 512), (512, 1100), (256, 768)):          # slabs: panel-aligned column ranges of S
        slab = _upload_native(s, col_begin=lo, col_end=hi)
        ys = _product(slab, a, k, c0[:, lo:hi].copy(), hi - lo, 1.3, 2.0).reshape(-1)
        assert_bits_equal(ys, full[lo:hi], f"slab [{lo},{hi}) against the whole matrix")
        slab.free()
    # y declared peer memory and beta == 1: the finish kernel adds at the owner (red.global.add), same bits
    L = sb.lib()
    x_d = torch.from_numpy(a).cuda()
    want = torch.from_numpy(c0.copy()).cuda()
    with _Env(SBLAS_B200_SPMV_NATIVE="1", SBLAS_B200_SPMV_DENSE="0"):
        whole.addmatmat_dev(x_d, 1, k, want, n, 1.3, 1.0)
        buf = torch.from_numpy(c0.copy()).cuda()
        assert L.sblas_b200_peer_register(C.c_void_p(buf.data_ptr()), buf.numel() * 4) == _capi.OK
        try:
            whole.addmatmat_dev(x_d, 1, k, buf, n, 1.3, 1.0)
            torch.cuda.synchronize()
        finally:
            assert L.sblas_b200_peer_unregister(C.c_void_p(buf.data_ptr())) == _capi.OK
    assert whole.last_product()["kernel"] == "spmv_native red", whole.last_product()["kernel"]
    assert_bits_equal(buf.cpu().numpy(), want.cpu().numpy(), "peer range")
    whole.free()


def test_native_stream_form_is_opt_in():
    # measured slower than the CSR row kernels on every shape (profiles/r02_spmv_native.md): never chosen on its own, and the
    # streams are only kept on the device when the upload is told to
    n = 8192
    tab = gen.codebook(255, 1)
    codes = gen.codes_bernoulli(n, n, 0.05, seed=17)
    S = sb.SparseMatrix(codes, n, n, n, tab, 255, sb.SparseMatrix.Trans)
    dev = S.device_matrix()
    a = gen.dense_uniform((1, n), 1); c0 = gen.dense_uniform((1, n), 2)
    dev.addmatmat_host(a, 1, n, c0.copy(), n, 1.0, 1.0)
    assert not dev.last_product()["kernel"].startswith("spmv_native")
    with _Env(SBLAS_B200_SPMV_NATIVE="1"):                      # asked for, but the streams were not kept: the usual kernels
        dev.addmatmat_host(a, 1, n, c0.copy(), n, 1.0, 1.0)
    assert not dev.last_product()["kernel"].startswith("spmv_native")
