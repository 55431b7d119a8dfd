"""The reference's own callers, compiled UNMODIFIED against this build, run on the GPU (SURVEY.md 8(b):
"unchanged call sites must work unmodified").

oracle/build_ref_callers.sh compiles, from /root/reference where they lie,
  src/sparse/sparse-matrix_test.cc:4-13, src/sparse/kernel_test.cc:27-57, src/test/blas_test.cc:31-47 (+ blas_test.h)
against sparsematrix_b200/src/sparse/*.h and lib/libsparsematrix.a into oracle/_ref/callers/ (built where the reference
is present, shipped to the GPU box with the snapshot).  Here they are only executed."""
import os
import re
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CALLERS = os.path.join(ROOT, "oracle", "_ref", "callers")


def _run(name, *args):
    exe = os.path.join(CALLERS, name)
    if not os.path.exists(exe):
        pytest.skip(f"{exe} not built (oracle/build_ref_callers.sh needs /root/reference)")
    env = dict(os.environ)
    libdir = os.path.join(CALLERS, "blas_libdir.txt")
    if os.path.exists(libdir):
        env["LD_LIBRARY_PATH"] = open(libdir).read().strip() + ":" + env.get("LD_LIBRARY_PATH", "")
    return subprocess.run([exe, *args], capture_output=True, text=True, timeout=600, env=env)


def test_reference_sparse_matrix_test_unmodified():
    r = _run("sparse-matrix_test")                       # SelfTest: KAT-1, KAT-2, round trip -- AddMatMat on the B200
    assert r.returncode == 0, r.stdout + r.stderr
    assert "SparseMatrix Test success" in r.stdout, r.stdout + r.stderr


def test_reference_kernel_test_unmodified():
    r = _run("kernel_test")                              # sblas_trans_kernel vs its naive transpose, 1023 x 511
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Kernel Test success" in r.stdout, r.stdout + r.stderr


def test_reference_blas_test_unmodified_sparse_row_on_the_gpu():
    # blas_test m n k check filter (blas_test.cc:31-37); its own check prints a message when > size/1e4 elements deviate
    # by more than 10 % from cblas_sgemm (blas_test.h:160-182) and is silent otherwise
    r = _run("blas_test", "128", "4096", "4096", "1", "sparse;cblas_sgemm_baseline")
    assert r.returncode == 0, r.stdout + r.stderr
    out = r.stdout + r.stderr
    assert "sgemm_sparse" in out, out                    # the row of its table
    assert not re.search(r"result failed", out), out     # sgemm_check's message (blas_test.h:168,176)
    r = _run("blas_test")                                # default sizes 117 x 1023 x 2047, all three rows
    assert r.returncode == 0, r.stdout + r.stderr
    assert "sgemm_sparse" in r.stdout + r.stderr
