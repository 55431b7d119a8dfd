"""Python face of the library: the reference's object interface and the raw device handle.

``SparseMatrix`` mirrors ``sblas::SparseMatrix<uint8,uint8,float>`` (reference
src/sparse/sparse-matrix.h:25-53): ``CopyForm / CopyTo / AddMatMat / NumRows / NumCols /
Destroy`` with the reference's argument order, backed by the C++ host class of this build
through ``include/sblas_b200_host.h``.  ``AddMatMat`` runs on the GPU.

``DeviceMatrix`` wraps the opaque ``sblas_b200_matrix`` of ``include/sblas_b200.h``:
uploads (whole matrix, panel-aligned slab, direct CSR), the host- and device-pointer
products, and the structure read-backs used by the parity tests.
"""
import ctypes as C
import os

import numpy as np

from . import _capi


def _ptr(x):
    """void* of a numpy array / torch tensor / int / None."""
    if x is None:
        return None
    if isinstance(x, int):
        return C.c_void_p(x)
    if isinstance(x, np.ndarray):
        return x.ctypes.data_as(C.c_void_p)
    if hasattr(x, "data_ptr"):
        return C.c_void_p(x.data_ptr())
    raise TypeError(type(x))


def _need(x, count, what):
    """Refuse buffers shorter than the library will touch (numpy arrays / torch tensors; raw ints are trusted)."""
    have = None
    if isinstance(x, np.ndarray):
        if x.dtype != np.float32 or not x.flags.c_contiguous:
            raise TypeError(f"{what}: expected a C-contiguous float32 array, got {x.dtype}"
                            f"{'' if x.flags.c_contiguous else ', not contiguous'}")
        have = x.size
    elif hasattr(x, "numel"):
        if str(getattr(x, "dtype", "torch.float32")) != "torch.float32" or not x.is_contiguous():
            raise TypeError(f"{what}: expected a contiguous float32 tensor, got {x.dtype}")
        have = x.numel()
    if have is not None and have < count:
        raise ValueError(f"{what} holds {have} elements, the call needs {count}")


class _PinnedBlock:
    """Owner of one sblas_b200_host_alloc() block; freed when the last array over it goes away."""

    def __init__(self, nbytes):
        p = C.c_void_p()
        _capi.check(_capi.lib().sblas_b200_host_alloc(nbytes, C.byref(p)))
        self.ptr, self.nbytes = p.value, nbytes

    def __del__(self):
        try:
            _capi.lib().sblas_b200_host_free(C.c_void_p(self.ptr))
        except Exception:
            pass


def host_buffer(shape, dtype=np.float32):
    """numpy array over page-locked, device-mapped host memory (``sblas_b200_host_alloc``): the kind of
    caller buffer for which ``AddMatMat`` / ``addmatmat_host`` overlap the copies with the product."""
    shape = (shape,) if isinstance(shape, int) else tuple(shape)
    nbytes = max(1, int(np.prod(shape)) * np.dtype(dtype).itemsize)
    block = _PinnedBlock(nbytes)
    raw = (C.c_uint8 * nbytes).from_address(block.ptr)
    raw._owner = block                                   # numpy keeps `raw` as the array's base, `raw` keeps the block
    return np.frombuffer(raw, dtype=dtype, count=int(np.prod(shape))).reshape(shape)


class DeviceMatrix:
    """W = S^T resident on one GPU (opaque sblas_b200_matrix)."""

    def __init__(self, handle, owned=True, owner=None):
        self._h = C.c_void_p(handle) if not isinstance(handle, C.c_void_p) else handle
        self._owned = owned
        self._owner = owner                 # a borrowed handle keeps the object that owns it alive

    # ---- construction -------------------------------------------------------------------
    @classmethod
    def upload(cls, pos, code, table, T, block_bounds, block_index_bounds, rows, cols, device=0,
               col_begin=None, col_end=None):
        """sblas_b200_upload / _upload_slab from the compressed members (numpy arrays)."""
        L = _capi.lib()
        pos = np.ascontiguousarray(pos, np.uint8); code = np.ascontiguousarray(code, np.uint8)
        table = np.ascontiguousarray(table, np.float32)
        bb = np.ascontiguousarray(block_bounds, np.int32).reshape(-1, 2)
        bib = np.ascontiguousarray(block_index_bounds, np.int32).reshape(-1, 2)
        out = C.c_void_p()
        if col_begin is None and col_end is None:
            rc = L.sblas_b200_upload(_ptr(pos), _ptr(code), pos.shape[0], _ptr(table), T, _ptr(bb), _ptr(bib),
                                     bb.shape[0], rows, cols, device, C.byref(out))
        else:
            rc = L.sblas_b200_upload_slab(_ptr(pos), _ptr(code), pos.shape[0], _ptr(table), T, _ptr(bb), _ptr(bib),
                                          bb.shape[0], rows, cols, int(col_begin or 0),
                                          int(cols if col_end is None else col_end), device, C.byref(out))
        _capi.check(rc)
        return cls(out)

    @classmethod
    def upload_csr(cls, row_ptr, col, code, table, T, n, k, device=0):
        L = _capi.lib()
        row_ptr = np.ascontiguousarray(row_ptr, np.int32); col = np.ascontiguousarray(col, np.int32)
        code = np.ascontiguousarray(code, np.uint8); table = np.ascontiguousarray(table, np.float32)
        out = C.c_void_p()
        _capi.check(L.sblas_b200_upload_csr(_ptr(row_ptr), _ptr(col), _ptr(code), _ptr(table), T, n, k, device,
                                            C.byref(out)))
        return cls(out)

    def free(self):
        if self._h and self._owned:
            _capi.lib().sblas_b200_free(self._h)
        self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    # ---- queries ------------------------------------------------------------------------
    def info(self):
        inf = _capi.Info()
        _capi.check(_capi.lib().sblas_b200_get_info(self._h, C.byref(inf)))
        return inf

    @property
    def n(self):
        return self.info().s_cols

    @property
    def k(self):
        return self.info().s_rows

    @property
    def nnz(self):
        return self.info().nnz

    def _shape(self):
        """(n, k), fetched once: a handle never changes shape."""
        sh = getattr(self, "_nk", None)
        if sh is None:
            inf = self.info()
            sh = self._nk = (inf.s_cols, inf.s_rows)
        return sh

    # ---- the product ----------------------------------------------------------------------
    def addmatmat_host(self, a, m, lda, c, ldc, alpha, beta):
        """C = beta*C + alpha*A*S on host buffers (numpy or pinned torch tensors), in place on c."""
        if m > 0:
            n, k = self._shape()
            _need(a, (m - 1) * lda + k, "a"); _need(c, (m - 1) * ldc + n, "c")
        _capi.check(_capi.lib().sblas_b200_addmatmat_host(self._h, _ptr(a), m, lda, _ptr(c), ldc, alpha, beta))
        return c

    def addmatmat_dev(self, d_a, m, lda, d_c, ldc, alpha, beta, stream=0):
        """Same on device pointers (torch CUDA tensors or raw ints), enqueued on `stream` (cudaStream_t as int)."""
        if m > 0:
            n, k = self._shape()
            _need(d_a, (m - 1) * lda + k, "d_a"); _need(d_c, (m - 1) * ldc + n, "d_c")
        _capi.check(_capi.lib().sblas_b200_addmatmat_dev(self._h, _ptr(d_a), m, lda, _ptr(d_c), ldc, alpha, beta,
                                                         C.c_void_p(stream) if stream else None))

    def addmatmat_dev_noalloc(self, d_a, m, lda, d_c, ldc, alpha, beta, stream=0):
        """sblas_b200_addmatmat_dev_noalloc: the same product without any allocation (fixed 4 MiB staging array,
        k * round_up(m, 4) <= 2^20); calls that overlap on one GPU race on that array."""
        if m > 0:
            n, k = self._shape()
            _need(d_a, (m - 1) * lda + k, "d_a"); _need(d_c, (m - 1) * ldc + n, "d_c")
        _capi.check(_capi.lib().sblas_b200_addmatmat_dev_noalloc(self._h, _ptr(d_a), m, lda, _ptr(d_c), ldc, alpha, beta,
                                                                 C.c_void_p(stream) if stream else None))

    def profile_begin(self):
        _capi.check(_capi.lib().sblas_b200_profile_begin(self._h))

    def profile_end(self):
        """-> (summed dominant-kernel ms, bracketed launches)"""
        ms = C.c_double(0.0); cnt = C.c_int64(0)
        _capi.check(_capi.lib().sblas_b200_profile_end(self._h, C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    def last_product(self):
        """sblas_b200_last_product: dominant kernel and algorithmic bytes / flops of the most recent product."""
        pi = _capi.ProductInfo()
        _capi.check(_capi.lib().sblas_b200_last_product(self._h, C.byref(pi)))
        return {"kernel": pi.kernel.decode(), "m": pi.m, "algorithmic_bytes": pi.algorithmic_bytes,
                "algorithmic_flops": pi.algorithmic_flops, "format_bytes_per_nnz": pi.format_bytes_per_nnz}

    # ---- structure read-back ----------------------------------------------------------------
    def decode_dense(self, stride, trans=False):
        inf = self.info()
        lead = inf.s_cols if trans else inf.s_rows
        dense = np.full((lead, stride), np.float32(1.0), np.float32)
        _capi.check(_capi.lib().sblas_b200_decode_dense(self._h, _ptr(dense), stride, int(bool(trans))))
        return dense

    def export_csr(self):
        inf = self.info()
        row_ptr = np.zeros(inf.s_cols + 1, np.int32)
        col = np.zeros(inf.nnz, np.int32); code = np.zeros(inf.nnz, np.uint8)
        _capi.check(_capi.lib().sblas_b200_export_csr(self._h, _ptr(row_ptr), _ptr(col), _ptr(code)))
        return row_ptr, col, code


def copyform_device(codes, rows, cols, stride, vals, T, trans=False, device=0, want_matrix=True):
    """sblas_b200_copyform_begin/_finish: CopyForm on the device (reference sparse-matrix.cc:21-99).
    codes: uint8 host array (rows x stride) or a torch CUDA tensor.
    -> (pos, code, block_bounds[nb,2], block_index_bounds[nb,2], DeviceMatrix or None)"""
    L = _capi.lib()
    if isinstance(codes, np.ndarray):
        codes = np.ascontiguousarray(codes, np.uint8).reshape(-1)
    vals = np.ascontiguousarray(vals, np.float32).reshape(-1)
    job = C.c_void_p(); entries = C.c_int64(0); nb = C.c_int32(0)
    _capi.check(L.sblas_b200_copyform_begin(_ptr(codes), rows, cols, stride, T, int(bool(trans)), device,
                                            C.byref(job), C.byref(entries), C.byref(nb)))
    pos = np.empty(entries.value, np.uint8); code = np.empty(entries.value, np.uint8)
    bb = np.zeros((nb.value, 2), np.int32); bib = np.zeros((nb.value, 2), np.int32)
    h = C.c_void_p()
    _capi.check(L.sblas_b200_copyform_finish(job, _ptr(pos), _ptr(code), _ptr(bb), _ptr(bib), _ptr(vals),
                                             C.byref(h) if want_matrix else None))
    return pos, code, bb, bib, (DeviceMatrix(h) if want_matrix else None)


def partition(block_bounds, block_index_bounds, cols, parts):
    """sblas_b200_partition: nnz-balanced panel-aligned column splits, length parts+1."""
    bb = np.ascontiguousarray(block_bounds, np.int32).reshape(-1, 2)
    bib = np.ascontiguousarray(block_index_bounds, np.int32).reshape(-1, 2)
    out = np.zeros(parts + 1, np.int32)
    _capi.check(_capi.lib().sblas_b200_partition(_ptr(bb), _ptr(bib), bb.shape[0], cols, parts, _ptr(out)))
    return out


def partition_csr(row_ptr, parts, align=256):
    """sblas_b200_partition_csr: nnz-balanced row splits of a CSR matrix, interior values multiples of `align`."""
    row_ptr = np.ascontiguousarray(row_ptr, np.int32)
    out = np.zeros(parts + 1, np.int32)
    _capi.check(_capi.lib().sblas_b200_partition_csr(_ptr(row_ptr), row_ptr.shape[0] - 1, parts, align, _ptr(out)))
    return out


def measure_fp32_peak(device=0):
    """sblas_b200_measure_fp32_peak: TFLOP/s of register-resident fp32 FMA chains on `device`."""
    t = C.c_double(0.0)
    _capi.check(_capi.lib().sblas_b200_measure_fp32_peak(device, C.byref(t)))
    return t.value


class SparseMatrix:
    """sblas::SparseMatrix<uint8,uint8,float> (host object of this build)."""

    NoTrans, Trans = 0, 1

    def __init__(self, density_matrix=None, rows=0, cols=0, stride=0, vals=None, val_table_size=0, trans=0):
        self._L = _capi.lib()
        self._h = C.c_void_p(self._L.sblas_host_create())
        if not self._h:
            raise MemoryError("sblas_host_create")
        if density_matrix is not None:
            self.CopyForm(density_matrix, rows, cols, stride, vals, val_table_size, trans)

    def __del__(self):
        try:
            if self._h:
                self._L.sblas_host_free(self._h)
                self._h = None
        except Exception:
            pass

    def Destroy(self):
        self._L.sblas_host_destroy(self._h)

    def CopyForm(self, density_matrix, rows, cols, stride, vals, val_table_size, trans=0):
        dm = np.ascontiguousarray(density_matrix, np.uint8).reshape(-1)
        if dm.shape[0] < (rows - 1) * stride + cols and rows > 0:
            raise ValueError("density_matrix smaller than rows x stride")
        vals = np.ascontiguousarray(vals, np.float32).reshape(-1)
        if vals.shape[0] < val_table_size:
            raise ValueError("vals shorter than val_table_size")
        rc = self._L.sblas_host_copyform(self._h, _ptr(dm), rows, cols, stride, _ptr(vals), val_table_size, int(trans))
        if rc != _capi.OK:
            raise ValueError("val_table_size outside [0, 255]")

    def CopyTo(self, stride, trans=0):
        """Returns the dense matrix (NumRows x stride, or NumCols x stride when trans)."""
        lead = self.NumCols() if trans else self.NumRows()
        minor = self.NumRows() if trans else self.NumCols()
        if stride < minor:
            raise ValueError(f"stride {stride} < {minor}: the dense output would be overrun")
        dense = np.full((lead, stride), np.float32(1.0), np.float32)
        self._L.sblas_host_copyto(self._h, _ptr(dense), stride, int(trans))
        return dense

    def AddMatMat(self, a, m, lda, c, ldc, alpha, beta):
        """In place on c; a and c are host buffers (numpy float32 / pinned torch tensors)."""
        if m > 0:
            _need(a, (m - 1) * lda + self.NumRows(), "a"); _need(c, (m - 1) * ldc + self.NumCols(), "c")
        _capi.check(self._L.sblas_host_addmatmat(self._h, _ptr(a), m, lda, _ptr(c), ldc, alpha, beta))
        return c

    def NumRows(self):
        return self._L.sblas_host_num_rows(self._h)

    def NumCols(self):
        return self._L.sblas_host_num_cols(self._h)

    def __eq__(self, other):
        return isinstance(other, SparseMatrix) and bool(self._L.sblas_host_equal(self._h, other._h))

    __hash__ = None

    # ---- B200 additions -----------------------------------------------------------------
    def members(self):
        """Copies of the compressed members: (pos, code, table[T+1], T, block_bounds[nb,2], block_index_bounds[nb,2])."""
        n = self._L.sblas_host_entries(self._h); nb = self._L.sblas_host_nblocks(self._h)
        T = self._L.sblas_host_table_size(self._h)
        p = [C.c_void_p() for _ in range(5)]
        self._L.sblas_host_members(self._h, *[C.byref(x) for x in p])

        def grab(ptr, count, dtype):
            if not ptr.value or count == 0:
                return np.zeros(count, dtype)
            buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr.value)
            return np.frombuffer(buf, dtype=dtype, count=count).copy()

        return (grab(p[0], n, np.uint8), grab(p[1], n, np.uint8), grab(p[2], T + 1 if T else 0, np.float32), T,
                grab(p[3], 2 * nb, np.int32).reshape(-1, 2), grab(p[4], 2 * nb, np.int32).reshape(-1, 2))

    def Save(self, path):
        """sblas_host_save: the compressed members to a file (see include/sblas_b200_host.h)."""
        _capi.check(self._L.sblas_host_save(self._h, os.fsencode(path)))

    def Load(self, path):
        """sblas_host_load: replace this object's contents with a saved matrix (validated first)."""
        _capi.check(self._L.sblas_host_load(self._h, os.fsencode(path)))

    def device_matrix(self):
        """The uploaded matrix of this object on the current CUDA device (borrowed: owned by this object, which the
        returned wrapper keeps alive; a later CopyForm / Destroy / load of this object still invalidates it)."""
        h = self._L.sblas_host_device_matrix(self._h)
        if not h:
            raise _capi.SblasError(_capi.ECUDA, self._L.sblas_b200_last_error().decode())
        return DeviceMatrix(h, owned=False, owner=self)

    def upload_slab(self, col_begin, col_end, device=0):
        pos, code, table, T, bb, bib = self.members()
        return DeviceMatrix.upload(pos, code, table, T, bb, bib, self.NumRows(), self.NumCols(), device, col_begin, col_end)


class SparseMatrix16:
    """sblas::SparseMatrix<uint16,uint8,float> of this build: two-byte position deltas (no fillers below gaps of 65 535).
    The reference has the code path (sparse-matrix.cc:24) but never instantiates it (:315)."""

    NoTrans, Trans = 0, 1

    def __init__(self, density_matrix=None, rows=0, cols=0, stride=0, vals=None, val_table_size=0, trans=0):
        self._L = _capi.lib()
        self._h = C.c_void_p(self._L.sblas_host16_create())
        if not self._h:
            raise MemoryError("sblas_host16_create")
        if density_matrix is not None:
            self.CopyForm(density_matrix, rows, cols, stride, vals, val_table_size, trans)

    def __del__(self):
        try:
            if self._h:
                self._L.sblas_host16_free(self._h)
                self._h = None
        except Exception:
            pass

    def CopyForm(self, density_matrix, rows, cols, stride, vals, val_table_size, trans=0):
        dm = np.ascontiguousarray(density_matrix, np.uint8).reshape(-1)
        if dm.shape[0] < (rows - 1) * stride + cols and rows > 0:
            raise ValueError("density_matrix smaller than rows x stride")
        vals = np.ascontiguousarray(vals, np.float32).reshape(-1)
        if vals.shape[0] < val_table_size:
            raise ValueError("vals shorter than val_table_size")
        if self._L.sblas_host16_copyform(self._h, _ptr(dm), rows, cols, stride, _ptr(vals), val_table_size, int(trans)) != _capi.OK:
            raise ValueError("val_table_size outside [0, 255]")

    def CopyTo(self, stride, trans=0):
        lead = self.NumCols() if trans else self.NumRows()
        minor = self.NumRows() if trans else self.NumCols()
        if stride < minor:
            raise ValueError(f"stride {stride} < {minor}: the dense output would be overrun")
        dense = np.full((lead, stride), np.float32(1.0), np.float32)
        self._L.sblas_host16_copyto(self._h, _ptr(dense), stride, int(trans))
        return dense

    def AddMatMat(self, a, m, lda, c, ldc, alpha, beta):
        if m > 0:
            _need(a, (m - 1) * lda + self.NumRows(), "a"); _need(c, (m - 1) * ldc + self.NumCols(), "c")
        _capi.check(self._L.sblas_host16_addmatmat(self._h, _ptr(a), m, lda, _ptr(c), ldc, alpha, beta))
        return c

    def NumRows(self):
        return self._L.sblas_host16_num_rows(self._h)

    def NumCols(self):
        return self._L.sblas_host16_num_cols(self._h)

    def members(self):
        """(pos uint16, code, table[T+1], T, block_bounds[nb,2], block_index_bounds[nb,2]) -- copies."""
        n = self._L.sblas_host16_entries(self._h); nb = self._L.sblas_host16_nblocks(self._h)
        T = self._L.sblas_host16_table_size(self._h)
        p = [C.c_void_p() for _ in range(5)]
        self._L.sblas_host16_members(self._h, *[C.byref(x) for x in p])

        def grab(ptr, count, dtype):
            if not ptr.value or count == 0:
                return np.zeros(count, dtype)
            buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr.value)
            return np.frombuffer(buf, dtype=dtype, count=count).copy()

        return (grab(p[0], n, np.uint16), grab(p[1], n, np.uint8), grab(p[2], T + 1 if T else 0, np.float32), T,
                grab(p[3], 2 * nb, np.int32).reshape(-1, 2), grab(p[4], 2 * nb, np.int32).reshape(-1, 2))

    def device_matrix(self):
        h = self._L.sblas_host16_device_matrix(self._h)
        if not h:
            raise _capi.SblasError(_capi.ECUDA, self._L.sblas_b200_last_error().decode())
        return DeviceMatrix(h, owned=False, owner=self)

