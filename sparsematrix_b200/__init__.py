"""sparsematrix_b200 -- B200 (sm_100a) implementation of the NeverLEX/sparsematrix
sparse x dense product path (SparseMatrix::AddMatMat: SpMV when m == 1, SpMM otherwise).

The product is CUDA only.  Importing this package does not load the shared library;
the first use does, and raises if it was not built (``__graft_entry__.build()``).
"""
from ._capi import LIB_PATH, SblasError, lib  # noqa: F401
from .matrix import (DeviceMatrix, SparseMatrix, SparseMatrix16, copyform_device, host_buffer, measure_fp32_peak, partition,  # noqa: F401
                     partition_csr)

__all__ = ["SparseMatrix", "SparseMatrix16", "DeviceMatrix", "partition", "partition_csr", "measure_fp32_peak", "copyform_device", "host_buffer", "SblasError", "lib", "LIB_PATH"]
