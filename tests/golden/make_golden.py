"""Generate tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref,
built from /root/reference by oracle/build_ref.sh) in this container.

    python tests/golden/make_golden.py

The GPU box has no /root/reference; these files are what travels.  Every case
stores its inputs (so no generator drift can masquerade as a parity failure),
the reference's compressed members after CopyForm, its CopyTo output in both
orientations and its AddMatMat result, all as exact bit patterns.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from sparsematrix_b200 import gen  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def run_case(name, codes, rows, cols, stride, vals, T, trans, a, m, lda, c, ldc, alpha, beta, copyto_strides):
    R = O.RefMatrix()
    R.CopyForm(codes, rows, cols, stride, vals, T, trans)
    s = R.stream()
    c_out = c.copy()
    R.AddMatMat(a.copy(), m, lda, c_out, ldc, alpha, beta)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        codes=codes, rows=rows, cols=cols, stride=stride, vals=vals, T=T, trans=int(trans),
        a=a, m=m, lda=lda, c_in=c, ldc=ldc, alpha=np.float32(alpha), beta=np.float32(beta),
        s_rows=s.rows, s_cols=s.cols, pos=s.pos, code=s.code, table=s.table,
        block_bounds=s.block_bounds, block_index_bounds=s.block_index_bounds,
        dense_notrans=R.CopyTo(copyto_strides[0], False), dense_trans=R.CopyTo(copyto_strides[1], True),
        stride_notrans=copyto_strides[0], stride_trans=copyto_strides[1],
        c_out=c_out)
    print(f"{name}: S {s.rows}x{s.cols} entries={s.entries} blocks={s.nblocks} m={m}")


def main():
    assert O.ref_selftest(), "reference SelfTest failed"
    f32 = np.float32
    tab8 = np.array([1.1, 2.2, 3.3, 4.4, 5.5, 6.6, 7.7, 8.8], f32)
    # KAT-1 / KAT-2: the inputs of SelfTest (sparse-matrix.cc:211-246)
    run_case("kat1", np.array([0, 255, 255, 3, 7, 255], np.uint8), 3, 2, 2, tab8, 8, False,
             np.array([3.1, 5, 7], f32), 1, 3, np.array([4, 8], f32), 2, 1.3, 2.0, (2, 3))
    run_case("kat2", np.array([0, 255, 7, 255, 3, 255], np.uint8), 2, 3, 3, tab8, 8, True,
             np.array([3.1, 5, 7], f32), 1, 3, np.array([4, 8], f32), 2, 1.3, 2.0, (2, 3))

    # one full panel + one narrow panel, NoTrans, small codebook, m not a multiple of anything
    rows, cols, stride = 70, 300, 304
    codes = np.full((rows, stride), 255, np.uint8)
    codes[:, :cols] = gen.codes_bernoulli(rows, cols, 0.2, seed=21, T=63)
    run_case("two_panels_notrans", codes, rows, cols, stride, gen.codebook(63, 5), 63, False,
             gen.dense_uniform((3, 80), 22), 3, 80, gen.dense_uniform((3, 310), 23), 310, 1.3, 2.0, (300, 70))

    # hypersparse, Trans orientation (the harness's), fillers dominate; lda/ldc padded
    n, k = 520, 1000
    codes = gen.codes_bernoulli(n, k, 0.004, seed=31)
    run_case("fillers_trans", codes, n, k, k, gen.codebook(255, 1), 255, True,
             gen.dense_uniform((9, 1003), 32), 9, 1003, gen.dense_uniform((9, 528), 33), 528, 1.8, 1.2, (520, 1000))

    # an all-zero middle panel (never recorded: sparse-matrix.cc:57) + beta-only columns; SpMV
    n, k = 700, 333
    codes = gen.codes_bernoulli(n, k, 0.05, seed=41)
    codes[256:512, :] = 255
    run_case("empty_panel_spmv", codes, n, k, k, gen.codebook(255, 1), 255, True,
             gen.dense_uniform((1, k), 42), 1, k, gen.dense_uniform((1, n), 43), n, 1.0, 0.5, (700, 333))

    # alpha == 0 (beta-scale only), beta == 0 (still multiplies), m = 8
    n, k = 260, 257
    codes = gen.codes_bernoulli(n, k, 0.3, seed=51)
    run_case("alpha0", codes, n, k, k, gen.codebook(255, 1), 255, True,
             gen.dense_uniform((8, k), 52), 8, k, gen.dense_uniform((8, n), 53), n, 0.0, 0.5, (260, 257))
    run_case("beta0_m8", codes, n, k, k, gen.codebook(255, 1), 255, True,
             gen.dense_uniform((8, k), 52), 8, k, gen.dense_uniform((8, n), 53), n, 1.0, 0.0, (260, 257))

    # the harness's default odd shape scaled down (117 x 1023 x 2047 -> 37 x 511 x 300), 25 % dense, alpha=beta=1
    n, k, m = 511, 300, 37
    codes = gen.codes_bernoulli(n, k, 0.25, seed=61)
    run_case("harness_like", codes, n, k, k, gen.codebook(255, 1), 255, True,
             gen.dense_uniform((m, k), 62), m, k, gen.dense_uniform((m, n), 63), n, 1.0, 1.0, (511, 300))


if __name__ == "__main__":
    main()
