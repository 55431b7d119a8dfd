"""ctypes binding of the C ABI (include/sblas_b200.h, include/sblas_b200_host.h).

The shared library is built in-tree by ``make -C sparsematrix_b200`` (``__graft_entry__.build()``)
as ``sparsematrix_b200/lib/libsblas_b200.so``.  There is no fallback of any kind: if the
library is missing, importing the symbols raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SBLAS_B200_LIB") or os.path.join(_HERE, "lib", "libsblas_b200.so")   # override: kernel experiments only

OK, EINVAL, ECUDA, ENOMEM, EFORMAT, EIO = 0, -1, -2, -3, -4, -5


class SblasError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"sblas_b200 error {code}: {msg}")
        self.code = code


class Info(C.Structure):
    _fields_ = [
        ("device", C.c_int32), ("s_rows", C.c_int32), ("s_cols", C.c_int32), ("col_begin", C.c_int32),
        ("table_size", C.c_int32), ("max_row_nnz", C.c_int32), ("long_rows", C.c_int32), ("spmv_lanes", C.c_int32),
        ("nnz", C.c_int64), ("fillers", C.c_int64), ("entries", C.c_int64), ("device_bytes", C.c_int64),
        ("row_len_hist", C.c_int64 * 33),
    ]


class ProductInfo(C.Structure):
    _fields_ = [("kernel", C.c_char * 64), ("m", C.c_int32), ("reserved", C.c_int32), ("algorithmic_bytes", C.c_int64),
                ("algorithmic_flops", C.c_int64), ("format_bytes_per_nnz", C.c_double)]


# name -> (restype, argtypes); every symbol the two headers declare
_vp, _i32, _i64, _f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
SYMBOLS = {
    # sblas_b200.h
    "sblas_b200_abi_version": (_i32, []),
    "sblas_b200_last_error": (C.c_char_p, []),
    "sblas_b200_device_count": (C.c_int, [_vp]),
    "sblas_b200_launch_count": (_i64, []),
    "sblas_b200_upload": (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "sblas_b200_upload_slab": (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "sblas_b200_upload_slab16": (C.c_int, [_vp, _vp, _i64, _vp, _i32, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp]),
    "sblas_b200_upload_csr": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "sblas_b200_partition": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp]),
    "sblas_b200_partition_csr": (C.c_int, [_vp, _i32, _i32, _i32, _vp]),
    "sblas_b200_last_product": (C.c_int, [_vp, _vp]),
    "sblas_b200_measure_fp32_peak": (C.c_int, [_i32, _vp]),
    "sblas_b200_free": (C.c_int, [_vp]),
    "sblas_b200_get_info": (C.c_int, [_vp, _vp]),
    "sblas_b200_addmatmat_host": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _i32, _f32, _f32]),
    "sblas_b200_addmatmat_dev_noalloc": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _i32, _f32, _f32, _vp]),
    "sblas_b200_addmatmat_dev": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _i32, _f32, _f32, _vp]),
    "sblas_b200_profile_begin": (C.c_int, [_vp]),
    "sblas_b200_profile_end": (C.c_int, [_vp, _vp, _vp]),
    "sblas_b200_decode_dense": (C.c_int, [_vp, _vp, _i32, _i32]),
    "sblas_b200_export_csr": (C.c_int, [_vp, _vp, _vp, _vp]),
    "sblas_b200_copyform_begin": (C.c_int, [_vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp, _vp]),
    "sblas_b200_copyform_finish": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "sblas_b200_copyform_abort": (None, [_vp]),
    "sblas_b200_scale_host": (C.c_int, [_vp, _i32, _i32, _i32, _f32]),
    "sblas_b200_transpose_host": (C.c_int, [_vp, _i32, _i32, _i32, _vp, _i32]),
    "sblas_b200_panel_product_host": (C.c_int, [_i32, _i32, _i32, _i32, _vp, _i32, _vp, _i32, _f32, _vp, _vp, _i32, _vp, _i32]),
    "sblas_b200_host_alloc": (C.c_int, [_i64, _vp]),
    "sblas_b200_host_free": (C.c_int, [_vp]),
    "sblas_b200_host_forget": (C.c_int, [_vp]),
    "sblas_b200_peer_alloc": (C.c_int, [_i32, _i64, _vp, _vp]),
    "sblas_b200_peer_open": (C.c_int, [_i32, _vp, _i64, _vp]),
    "sblas_b200_peer_close": (C.c_int, [_vp]),
    "sblas_b200_peer_free": (C.c_int, [_vp]),
    "sblas_b200_peer_register": (C.c_int, [_vp, _i64]),
    "sblas_b200_peer_unregister": (C.c_int, [_vp]),
    # sblas_b200_host.h
    "sblas_host_create": (_vp, []),
    "sblas_host_free": (None, [_vp]),
    "sblas_host_destroy": (None, [_vp]),
    "sblas_host_copyform": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _i32, _i32]),
    "sblas_host_copyto": (None, [_vp, _vp, _i32, _i32]),
    "sblas_host_addmatmat": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _i32, _f32, _f32]),
    "sblas_host_num_rows": (_i32, [_vp]),
    "sblas_host_num_cols": (_i32, [_vp]),
    "sblas_host_equal": (_i32, [_vp, _vp]),
    "sblas_host_selftest": (_i32, []),
    "sblas_host_entries": (_i64, [_vp]),
    "sblas_host_nblocks": (_i32, [_vp]),
    "sblas_host_table_size": (_i32, [_vp]),
    "sblas_host_members": (None, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "sblas_host_device_matrix": (_vp, [_vp]),
    "sblas_host16_create": (_vp, []),
    "sblas_host16_free": (None, [_vp]),
    "sblas_host16_copyform": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _i32, _i32]),
    "sblas_host16_copyto": (None, [_vp, _vp, _i32, _i32]),
    "sblas_host16_addmatmat": (C.c_int, [_vp, _vp, _i32, _i32, _vp, _i32, _f32, _f32]),
    "sblas_host16_num_rows": (_i32, [_vp]),
    "sblas_host16_num_cols": (_i32, [_vp]),
    "sblas_host16_entries": (_i64, [_vp]),
    "sblas_host16_nblocks": (_i32, [_vp]),
    "sblas_host16_table_size": (_i32, [_vp]),
    "sblas_host16_members": (None, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "sblas_host16_device_matrix": (_vp, [_vp]),
    "sblas_host_save": (C.c_int, [_vp, C.c_char_p]),
    "sblas_host_load": (C.c_int, [_vp, C.c_char_p]),
}

_lib = None


def lib():
    """The loaded library with typed entry points.  Raises if it was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C {_HERE}` "
                              "(or __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)          # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc != OK:
        raise SblasError(rc, lib().sblas_b200_last_error().decode("utf-8", "replace"))
