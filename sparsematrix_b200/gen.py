"""Deterministic synthetic inputs for the parity tests, smoke() and bench.py.

The reference's harness draws its inputs with unseeded ``rand()`` / ``srand(time)``
(src/test/blas_test.h:119-147, blas_test.cc:39), so nothing it generates is
reproducible; these generators keep its *shape* (a dense uint8 code matrix
``n x k`` handed to ``CopyForm(..., SblasTrans)``, code 255 = "zero", a
255-entry fp32 codebook -- blas_test.h:132-147) but take explicit seeds
(SURVEY.md 8(d)).  numpy only; no device work here.
"""
import numpy as np

ZERO_CODE = 255  # T <= 255, so 255 is never a valid code (sparse-matrix.cc:25,44)


def codebook(T=255, seed=1):
    """T fp32 codebook entries ~ U(-1,1), none exactly 0 (a stored entry is a non-zero)."""
    rng = np.random.default_rng(seed)
    tab = rng.uniform(-1.0, 1.0, size=T).astype(np.float32)
    tab[tab == 0] = np.float32(0.5)
    return tab


def codes_bernoulli(n, k, density, seed, T=255, chunk_rows=1024):
    """Dense ``n x k`` uint8 code matrix: each element is a non-zero with probability
    ``density`` (code ~ U{0..T-1}), else ZERO_CODE.  Row j is output row j of W = S^T."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, k), np.uint8)
    for r0 in range(0, n, chunk_rows):
        r1 = min(n, r0 + chunk_rows)
        keep = rng.random((r1 - r0, k), dtype=np.float32) < np.float32(density)
        vals = rng.integers(0, T, size=(r1 - r0, k), dtype=np.uint8)
        out[r0:r1] = np.where(keep, vals, np.uint8(ZERO_CODE))
    return out


def dense_uniform(shape, seed):
    rng = np.random.default_rng(seed)
    return rng.uniform(-1.0, 1.0, size=shape).astype(np.float32)


def powerlaw_row_lengths(n, seed, x_min=8, cap=1 << 20):
    """Pareto(alpha=2, x_min) row lengths, clamp [1, cap]: L = floor(x_min * U^(-1/2)); mean ~ 2*x_min."""
    rng = np.random.default_rng(seed)
    u = rng.random(n)
    u = np.maximum(u, 1e-300)
    L = np.floor(x_min / np.sqrt(u))
    return np.clip(L, 1, cap).astype(np.int64)


def powerlaw_csr(n, k, seed, x_min=8, cap=1 << 20, T=255, chunk_rows=1 << 18):
    """CSR(W) with skewed row lengths for shapes the panel format cannot hold
    (BASELINE.json config 4).  Inside a row the columns are sorted and distinct:
    jittered strata  col_t = floor((t + u_t) * k / L),  u_t ~ U[0,1).
    Returns (row_ptr int32[n+1], col int32[nnz], code uint8[nnz])."""
    lens = np.minimum(powerlaw_row_lengths(n, seed, x_min, cap), k)
    row_ptr64 = np.zeros(n + 1, np.int64)
    np.cumsum(lens, out=row_ptr64[1:])
    nnz = int(row_ptr64[-1])
    if nnz >= 2 ** 31:
        raise ValueError("nnz does not fit int32")
    col = np.empty(nnz, np.int32)
    code = np.empty(nnz, np.uint8)
    rng = np.random.default_rng(seed + 1)
    for r0 in range(0, n, chunk_rows):
        r1 = min(n, r0 + chunk_rows)
        e0, e1 = int(row_ptr64[r0]), int(row_ptr64[r1])
        ln = lens[r0:r1]
        rep_len = np.repeat(ln, ln)
        t = np.arange(e1 - e0, dtype=np.int64) - np.repeat(row_ptr64[r0:r1] - e0, ln)
        lo = (t * k) // rep_len                      # stratum t = [lo, hi): disjoint, non-empty since L <= k
        hi = ((t + 1) * k) // rep_len
        c = lo + np.floor(rng.random(e1 - e0) * (hi - lo)).astype(np.int64)
        np.minimum(c, hi - 1, out=c)
        col[e0:e1] = c
        code[e0:e1] = rng.integers(0, T, size=e1 - e0, dtype=np.uint8)
    return row_ptr64.astype(np.int32), col, code
