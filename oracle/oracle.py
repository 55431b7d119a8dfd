"""ctypes front-end for the CPU checker.  TEST INFRASTRUCTURE -- see sblas_oracle.c.

Two libraries:
  * ``liboracle.so``  -- the plain-C restatement (oracle/sblas_oracle.c)
  * ``_ref/libsparsematrix_ref.so`` -- the unmodified reference, compiled from
    /root/reference by oracle/build_ref.sh (class ``RefMatrix``); present in
    this container and shipped prebuilt to the GPU box.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product (sparsematrix_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(_HERE, "liboracle.so")
_REF_SO = os.path.join(_HERE, "_ref", "libsparsematrix_ref.so")

_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def build(force=False):
    """Compile liboracle.so (always possible: gcc) and _ref (when /root/reference exists)."""
    src = os.path.join(_HERE, "sblas_oracle.c")
    if force or not os.path.exists(_ORACLE_SO) or os.path.getmtime(_ORACLE_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    ref_src = os.environ.get("SBLAS_REFERENCE_ROOT", "/root/reference")
    if os.path.isdir(ref_src) and (force or not os.path.exists(_REF_SO)):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_ORACLE_SO)
        vp = C.c_void_p
        i32, i64, f32 = C.c_int32, C.c_int64, C.c_float
        L.sbo_copyform.restype = i64
        L.sbo_copyform.argtypes = [vp, i32, i32, i32, vp, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
        L.sbo_copyto.restype = None
        L.sbo_copyto.argtypes = [vp, vp, vp, i32, vp, vp, i32, i32, i32, vp, i32, i32]
        L.sbo_copyform16.restype = i64
        L.sbo_copyform16.argtypes = L.sbo_copyform.argtypes
        L.sbo_copyto16.restype = None
        L.sbo_copyto16.argtypes = L.sbo_copyto.argtypes
        L.sbo_stream_to_csr16.restype = i64
        L.sbo_stream_to_csr16.argtypes = [vp, vp, i32, vp, vp, i32, i32, i32, vp, vp, vp, vp]
        L.sbo_addmatmat.restype = None
        L.sbo_addmatmat.argtypes = [vp, vp, vp, i32, vp, vp, i32, i32, i32, vp, i32, i32, vp, i32, f32, f32]
        L.sbo_addmatmat_f64.restype = None
        L.sbo_addmatmat_f64.argtypes = [vp, vp, vp, i32, vp, vp, i32, i32, i32, vp, i32, i32, vp, i32, f32, f32, vp, vp]
        L.sbo_stream_to_csr.restype = i64
        L.sbo_stream_to_csr.argtypes = [vp, vp, i32, vp, vp, i32, i32, i32, vp, vp, vp, vp]
        L.sbo_addmatmat_csr.restype = None
        L.sbo_addmatmat_csr.argtypes = [vp, vp, vp, vp, i32, i32, i32, vp, i32, i32, vp, i32, f32, f32]
        L.sbo_addmatmat_csr_omp.restype = i32
        L.sbo_addmatmat_csr_omp.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp, f32, f32]
        L.sbo_addmatmat_csr_f64.restype = None
        L.sbo_addmatmat_csr_f64.argtypes = [vp, vp, vp, vp, i32, i32, i32, vp, i32, i32, vp, i32, f32, f32, vp, vp]
        L.sbo_transpose.restype = None
        L.sbo_transpose.argtypes = [vp, i32, i32, i32, vp, i32]
        L.sbo_error_measures.restype = None
        L.sbo_error_measures.argtypes = [vp, i32, i32, i32, vp, vp, vp]
        _lib = L
    return _lib


def have_ref():
    if not os.path.exists(_REF_SO):
        try:
            build()
        except Exception:
            pass
    return os.path.exists(_REF_SO)


def ref():
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError("oracle/_ref/libsparsematrix_ref.so is missing (run oracle/build_ref.sh where /root/reference exists)")
        R = C.CDLL(_REF_SO)
        vp = C.c_void_p
        i32, i64, f32 = C.c_int32, C.c_int64, C.c_float
        R.sref_create.restype = vp
        R.sref_free.argtypes = [vp]
        R.sref_copyform.argtypes = [vp, vp, i32, i32, i32, vp, i32, i32]
        R.sref_copyto.argtypes = [vp, vp, i32, i32]
        R.sref_addmatmat.argtypes = [vp, vp, i32, i32, vp, i32, f32, f32]
        R.sref_rows.argtypes = [vp]; R.sref_rows.restype = i32
        R.sref_cols.argtypes = [vp]; R.sref_cols.restype = i32
        R.sref_entries.argtypes = [vp]; R.sref_entries.restype = i64
        R.sref_nblocks.argtypes = [vp]; R.sref_nblocks.restype = i32
        R.sref_table_len.argtypes = [vp]; R.sref_table_len.restype = i32
        R.sref_export.argtypes = [vp, vp, vp, vp, vp, vp]
        R.sref_selftest.restype = i32
        R.sref_trans_kernel.argtypes = [vp, i32, i32, i32, vp, i32]
        R.sref_beta_kernel.argtypes = [vp, i32, i32, i32, f32]
        R.sref_panel_kernel.argtypes = [i32, i32, i32, i32, vp, i32, vp, i32, f32, vp, vp, i32, vp, i32]
        _ref = R
    return _ref


def _p(arr):
    return arr.ctypes.data_as(C.c_void_p) if arr is not None else None


class Stream:
    """The compressed members of sblas::SparseMatrix (sparse-matrix.h:45-52) as numpy arrays."""

    __slots__ = ("pos", "code", "table", "T", "block_bounds", "block_index_bounds", "rows", "cols")

    def __init__(self, pos, code, table, T, block_bounds, block_index_bounds, rows, cols):
        self.pos, self.code, self.table, self.T = pos, code, table, int(T)
        self.block_bounds, self.block_index_bounds = block_bounds, block_index_bounds
        self.rows, self.cols = int(rows), int(cols)

    @property
    def entries(self):
        return int(self.pos.shape[0])

    @property
    def nblocks(self):
        return int(self.block_bounds.shape[0])

    def same_as(self, other):
        return (self.rows == other.rows and self.cols == other.cols and self.T == other.T
                and np.array_equal(self.pos, other.pos) and np.array_equal(self.code, other.code)
                and np.array_equal(self.table.view(np.uint32), other.table.view(np.uint32))
                and np.array_equal(self.block_bounds, other.block_bounds)
                and np.array_equal(self.block_index_bounds, other.block_index_bounds))


# --------------------------------------------------------------------------- C restatement

def copyform(codes, rows, cols, stride, vals, T, trans=False, wide=False):
    """oracle CopyForm -> Stream.  codes: uint8 array holding >= rows*stride bytes.
    wide: two-byte position deltas, the SparseMatrix<uint16, uint8, float> instantiation (Stream.pos is uint16)."""
    L = lib()
    fn = L.sbo_copyform16 if wide else L.sbo_copyform
    codes = np.ascontiguousarray(codes, dtype=np.uint8).reshape(-1)
    vals = np.ascontiguousarray(vals, dtype=np.float32).reshape(-1)
    nb = C.c_int32(0); sr = C.c_int32(0); sc = C.c_int32(0)
    n = fn(_p(codes), rows, cols, stride, _p(vals), T, int(bool(trans)),
           None, None, None, None, None, C.byref(nb), C.byref(sr), C.byref(sc))
    if n < 0:
        raise ValueError("val_table_size out of range")
    pos = np.empty(n, np.uint16 if wide else np.uint8); code = np.empty(n, np.uint8)
    table = np.zeros(T + 1 if T > 0 else 0, np.float32)
    bb = np.zeros((nb.value, 2), np.int32); bib = np.zeros((nb.value, 2), np.int32)
    fn(_p(codes), rows, cols, stride, _p(vals), T, int(bool(trans)),
       _p(pos), _p(code), _p(table) if T > 0 else None, _p(bb), _p(bib),
       C.byref(nb), C.byref(sr), C.byref(sc))
    return Stream(pos, code, table, T, bb, bib, sr.value, sc.value)


def copyto(s, stride, trans=False):
    lead = s.cols if trans else s.rows
    dense = np.full((lead, stride), np.float32(1.0), np.float32)  # reference memsets; prove it
    fn = lib().sbo_copyto16 if s.pos.dtype == np.uint16 else lib().sbo_copyto
    fn(_p(s.pos), _p(s.code), _p(s.table), s.T, _p(s.block_bounds), _p(s.block_index_bounds),
       s.nblocks, s.rows, s.cols, _p(dense), stride, int(bool(trans)))
    return dense


def addmatmat(s, a, m, lda, c, ldc, alpha, beta):
    """In place on c (float32, C-contiguous, >= m*ldc)."""
    assert a.dtype == np.float32 and c.dtype == np.float32
    lib().sbo_addmatmat(_p(s.pos), _p(s.code), _p(s.table), s.T, _p(s.block_bounds), _p(s.block_index_bounds),
                        s.nblocks, s.rows, s.cols, _p(a), m, lda, _p(c), ldc, alpha, beta)
    return c


def addmatmat_f64(s, a, m, lda, c, ldc, alpha, beta):
    y = np.empty((m, s.cols), np.float64); bd = np.empty((m, s.cols), np.float64)
    lib().sbo_addmatmat_f64(_p(s.pos), _p(s.code), _p(s.table), s.T, _p(s.block_bounds), _p(s.block_index_bounds),
                            s.nblocks, s.rows, s.cols, _p(a), m, lda, _p(c), ldc, alpha, beta, _p(y), _p(bd))
    return y, bd


def stream_to_csr(s):
    """-> (row_ptr[n+1] int32, col[nnz] int32, code[nnz] uint8, fillers)"""
    L = lib()
    fill = C.c_int64(0)
    row_ptr = np.zeros(s.cols + 1, np.int32)
    fn = L.sbo_stream_to_csr16 if s.pos.dtype == np.uint16 else L.sbo_stream_to_csr
    nnz = fn(_p(s.pos), _p(s.code), s.T, _p(s.block_bounds), _p(s.block_index_bounds), s.nblocks,
             s.rows, s.cols, _p(row_ptr), None, None, C.byref(fill))
    col = np.empty(nnz, np.int32); code = np.empty(nnz, np.uint8)
    fn(_p(s.pos), _p(s.code), s.T, _p(s.block_bounds), _p(s.block_index_bounds), s.nblocks,
       s.rows, s.cols, _p(row_ptr), _p(col), _p(code), C.byref(fill))
    return row_ptr, col, code, fill.value


def addmatmat_csr(row_ptr, col, code, table, T, n, k, a, m, lda, c, ldc, alpha, beta):
    lib().sbo_addmatmat_csr(_p(row_ptr), _p(col), _p(code), _p(table), T, n, k, _p(a), m, lda, _p(c), ldc, alpha, beta)
    return c


def addmatmat_csr_omp(row_ptr, col, code, table, T, n, x, y, alpha, beta):
    """m == 1 on all host cores (OpenMP over rows; same bits as addmatmat_csr).  Returns the thread count used."""
    return int(lib().sbo_addmatmat_csr_omp(_p(row_ptr), _p(col), _p(code), _p(table), T, n, _p(x), _p(y), alpha, beta))


def addmatmat_csr_f64(row_ptr, col, code, table, T, n, k, a, m, lda, c, ldc, alpha, beta):
    y = np.empty((m, n), np.float64); bd = np.empty((m, n), np.float64)
    lib().sbo_addmatmat_csr_f64(_p(row_ptr), _p(col), _p(code), _p(table), T, n, k, _p(a), m, lda, _p(c), ldc,
                                alpha, beta, _p(y), _p(bd))
    return y, bd


def transpose(a, m, n, lda, ldsa):
    sa = np.zeros((n, ldsa), np.float32)
    lib().sbo_transpose(_p(a), m, n, lda, _p(sa), ldsa)
    return sa


def error_measures(y, m, n, ldy, y64, bound):
    """-> (backward, normwise, elementwise) max errors of fp32 y against the fp64 statement."""
    out = np.zeros(3, np.float64)
    lib().sbo_error_measures(_p(y), m, n, ldy, _p(y64), _p(bound), _p(out))
    return float(out[0]), float(out[1]), float(out[2])


# --------------------------------------------------------------------------- the real reference

class RefMatrix:
    """sblas::SparseMatrix<uint8,uint8,float> of the unmodified reference (oracle/_ref)."""

    def __init__(self):
        self._R = ref()
        self._h = C.c_void_p(self._R.sref_create())

    def __del__(self):
        try:
            if self._h:
                self._R.sref_free(self._h)
                self._h = None
        except Exception:
            pass

    def CopyForm(self, codes, rows, cols, stride, vals, T, trans=False):
        codes = np.ascontiguousarray(codes, dtype=np.uint8).reshape(-1)
        vals = np.ascontiguousarray(vals, dtype=np.float32).reshape(-1)
        self._R.sref_copyform(self._h, _p(codes), rows, cols, stride, _p(vals), T, int(bool(trans)))

    def CopyTo(self, stride, trans=False):
        lead = self.NumCols() if trans else self.NumRows()
        dense = np.full((lead, stride), np.float32(1.0), np.float32)
        self._R.sref_copyto(self._h, _p(dense), stride, int(bool(trans)))
        return dense

    def AddMatMat(self, a, m, lda, c, ldc, alpha, beta):
        assert a.dtype == np.float32 and c.dtype == np.float32
        self._R.sref_addmatmat(self._h, _p(a), m, lda, _p(c), ldc, alpha, beta)
        return c

    def NumRows(self):
        return self._R.sref_rows(self._h)

    def NumCols(self):
        return self._R.sref_cols(self._h)

    def stream(self):
        n = self._R.sref_entries(self._h); nb = self._R.sref_nblocks(self._h); tl = self._R.sref_table_len(self._h)
        pos = np.empty(n, np.uint8); code = np.empty(n, np.uint8); table = np.zeros(tl, np.float32)
        bb = np.zeros((nb, 2), np.int32); bib = np.zeros((nb, 2), np.int32)
        self._R.sref_export(self._h, _p(pos), _p(code), _p(table), _p(bb), _p(bib))
        return Stream(pos, code, table, max(tl - 1, 0), bb, bib, self.NumRows(), self.NumCols())


_REF16_SO = os.path.join(_HERE, "_ref", "libsparsematrix_ref16.so")
_ref16 = None


def have_ref16():
    return os.path.exists(_REF16_SO)


def ref16():
    global _ref16
    if _ref16 is None:
        if not have_ref16():
            raise RuntimeError("oracle/_ref/libsparsematrix_ref16.so is missing (oracle/build_ref.sh builds it where /root/reference exists)")
        R = C.CDLL(_REF16_SO)
        vp = C.c_void_p
        i32, i64, f32 = C.c_int32, C.c_int64, C.c_float
        R.sref16_create.restype = vp
        R.sref16_free.argtypes = [vp]
        R.sref16_copyform.argtypes = [vp, vp, i32, i32, i32, vp, i32, i32]
        R.sref16_copyto.argtypes = [vp, vp, i32, i32]
        R.sref16_addmatmat.argtypes = [vp, vp, i32, i32, vp, i32, f32, f32]
        for name, res in (("rows", i32), ("cols", i32), ("entries", i64), ("nblocks", i32), ("table_len", i32)):
            fn = getattr(R, "sref16_" + name)
            fn.argtypes = [vp]; fn.restype = res
        R.sref16_export.argtypes = [vp, vp, vp, vp, vp, vp]
        _ref16 = R
    return _ref16


class Ref16Matrix:
    """sblas::SparseMatrix<uint16,uint8,float> of the unmodified reference (oracle/ref_shim16.cc): two-byte position deltas."""

    def __init__(self):
        self._R = ref16()
        self._h = C.c_void_p(self._R.sref16_create())

    def __del__(self):
        try:
            if self._h:
                self._R.sref16_free(self._h)
                self._h = None
        except Exception:
            pass

    def CopyForm(self, codes, rows, cols, stride, vals, T, trans=False):
        codes = np.ascontiguousarray(codes, dtype=np.uint8).reshape(-1)
        vals = np.ascontiguousarray(vals, dtype=np.float32).reshape(-1)
        self._R.sref16_copyform(self._h, _p(codes), rows, cols, stride, _p(vals), T, int(bool(trans)))

    def CopyTo(self, stride, trans=False):
        lead = self.NumCols() if trans else self.NumRows()
        dense = np.full((lead, stride), np.float32(1.0), np.float32)
        self._R.sref16_copyto(self._h, _p(dense), stride, int(bool(trans)))
        return dense

    def AddMatMat(self, a, m, lda, c, ldc, alpha, beta):
        assert a.dtype == np.float32 and c.dtype == np.float32
        self._R.sref16_addmatmat(self._h, _p(a), m, lda, _p(c), ldc, alpha, beta)
        return c

    def NumRows(self):
        return self._R.sref16_rows(self._h)

    def NumCols(self):
        return self._R.sref16_cols(self._h)

    def stream(self):
        n = self._R.sref16_entries(self._h); nb = self._R.sref16_nblocks(self._h); tl = self._R.sref16_table_len(self._h)
        pos = np.empty(n, np.uint16); code = np.empty(n, np.uint8); table = np.zeros(tl, np.float32)
        bb = np.zeros((nb, 2), np.int32); bib = np.zeros((nb, 2), np.int32)
        self._R.sref16_export(self._h, _p(pos), _p(code), _p(table), _p(bb), _p(bib))
        return Stream(pos, code, table, max(tl - 1, 0), bb, bib, self.NumRows(), self.NumCols())


def ref_selftest():
    return bool(ref().sref_selftest())


def ref_trans_kernel(a, m, n, lda, ldsa):
    sa = np.zeros((n, ldsa), np.float32)
    ref().sref_trans_kernel(_p(a), m, n, lda, _p(sa), ldsa)
    return sa


def ref_panel_kernel(which, m, n, k, a, lda, c, ldc, alpha, ppos, pval, table, T):
    """Run one of the reference's four panel formulations (kernel.cc:213-369, :771-800) in place on c."""
    ref().sref_panel_kernel(which, m, n, k, _p(a), lda, _p(c), ldc, alpha, _p(ppos), _p(pval), int(ppos.shape[0]), _p(table), T)
    return c
