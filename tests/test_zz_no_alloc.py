"""sblas_b200_addmatmat_dev_noalloc: the product with its staging in a fixed device array (no allocation).

Collected last on purpose: a failure here hides nothing after it.
The new entry point is called through the C ABI and compared with the existing device entry point,
sblas_b200_addmatmat_dev, on the same inputs -- two calls in a row (the array keeps the first call's
contents) -- within the tolerance of the parity suite: backward error
  |c_new - c_old| / (|beta c| + |alpha| sum |a||s|) <= 1e-5.
"""
import numpy as np
import pytest

import sparsematrix_b200 as sb
from sparsematrix_b200 import _capi, gen
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL = 1e-5


def _both(dev, s, m, alpha, beta, seed, torch):
    n, k = s.cols, s.rows
    a = gen.dense_uniform((m, k), seed)
    c0 = gen.dense_uniform((m, n), seed + 1)
    d_a = torch.from_numpy(a).cuda()
    d_old = torch.from_numpy(c0).cuda()
    d_new = torch.from_numpy(c0).cuda()
    dev.addmatmat_dev(d_a, m, k, d_old, n, alpha, beta)
    dev.addmatmat_dev_noalloc(d_a, m, k, d_new, n, alpha, beta)
    torch.cuda.synchronize()
    old, new = d_old.cpu().numpy(), d_new.cpu().numpy()
    y64, bound = O.addmatmat_f64(s, a, m, k, c0, n, alpha, beta)
    diff = np.abs(new.astype(np.float64) - old.astype(np.float64))
    with np.errstate(divide="ignore", invalid="ignore"):
        q = np.where(bound > 0, diff / bound, np.where(diff == 0, 0.0, np.inf))
    assert q.max() <= TOL, f"m={m} alpha={alpha} beta={beta}: new vs existing entry point differ by {q.max():.3e} (backward)"
    # and both against the fp64 statement, like every parity test
    for got, what in ((old, "existing"), (new, "noalloc")):
        back, _, _ = O.error_measures(got, m, n, n, y64, bound)
        assert back <= TOL, f"{what} entry point, m={m}: backward error {back:.3e}"


def test_noalloc_entry_point_matches_the_existing_one_on_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    n, k, T = 1000, 777, 255                       # ragged last panel, k not a multiple of 32
    codes = gen.codes_bernoulli(n, k, 0.07, seed=31)
    tab = gen.codebook(T, 1)
    s = O.copyform(codes, n, k, k, tab, T, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    launches0 = sb.lib().sblas_b200_launch_count()
    # first call: wide m (padding columns of the staging array are written as zeros);
    # second call: a narrower m with other scalars, over the first call's left-overs
    _both(dev, s, 37, 1.3, 2.0, 5, torch)
    _both(dev, s, 5, 1.0, 1.0, 9, torch)
    _both(dev, s, 1, 1.8, 1.2, 13, torch)          # m = 1 through the same path
    assert sb.lib().sblas_b200_launch_count() > launches0
    dev.free()


def test_noalloc_entry_point_rejects_sizes_beyond_its_fixed_scratch():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    n, k, T = 512, 4096, 255
    codes = gen.codes_bernoulli(n, k, 0.01, seed=32)
    s = O.copyform(codes, n, k, k, gen.codebook(T, 1), T, True)
    dev = sb.DeviceMatrix.upload(s.pos, s.code, s.table, s.T, s.block_bounds, s.block_index_bounds, s.rows, s.cols)
    m = 260                                        # k * round_up(m, 4) = 4096 * 260 > 2^20
    d_a = torch.zeros((m, k), device="cuda")
    d_c = torch.ones((m, n), device="cuda")
    with pytest.raises(_capi.SblasError) as err:
        dev.addmatmat_dev_noalloc(d_a, m, k, d_c, n, 1.0, 1.0)
    assert "scratch" in str(err.value) and "SBLAS_B200_NOALLOC_MAX_FLOATS" in str(err.value)
    torch.cuda.synchronize()
    assert bool((d_c == 1).all())                  # nothing was enqueued
    dev.addmatmat_dev_noalloc(d_a, 256, k, d_c, n, 1.0, 0.5)     # exactly at the limit: accepted (A = 0: C is scaled)
    torch.cuda.synchronize()
    assert bool((d_c[:256] == 0.5).all()) and bool((d_c[256:] == 1).all())
    dev.free()
