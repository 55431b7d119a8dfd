#!/usr/bin/env python
"""m = 1 product of an n x n Bernoulli matrix: the native-stream kernel (SBLAS_B200_SPMV_NATIVE=1) against what the library
chooses without it (=0), cold (L2 flushed) and warm.  Usage: tools/spmv_native_probe.py n:density [n:density ...]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparsematrix_b200 as sb  # noqa: E402
from sparsematrix_b200 import gen  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
tab = gen.codebook()
os.environ["SBLAS_B200_KEEP_NATIVE"] = "1"
for spec in sys.argv[1:]:
    n, dens = spec.split(":")
    n, dens = int(n), float(dens)
    codes = gen.codes_bernoulli(n, n, dens, 12)
    S = sb.SparseMatrix(codes, n, n, n, tab, 255, sb.SparseMatrix.Trans)
    mat = S.device_matrix()
    nnz = int((codes != 255).sum())
    x = torch.from_numpy(gen.dense_uniform((1, n), 2)).cuda()
    ref = None
    for mode in ("0", "1"):
        os.environ["SBLAS_B200_SPMV_NATIVE"] = mode
        y = torch.zeros((1, n), device="cuda")
        for _ in range(3):
            mat.addmatmat_dev(x, 1, n, y, n, 1.0, 1.0)
        evs = []
        for _ in range(20):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); mat.addmatmat_dev(x, 1, n, y, n, 1.0, 1.0); b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        cold = np.array([a.elapsed_time(b) for a, b in evs]) * 1e3
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20):
            mat.addmatmat_dev(x, 1, n, y, n, 1.0, 1.0)
        b.record(); torch.cuda.synchronize()
        warm = a.elapsed_time(b) / 20 * 1e3
        y.zero_(); mat.addmatmat_dev(x, 1, n, y, n, 1.0, 1.0); torch.cuda.synchronize()
        got = y.cpu().numpy()
        if ref is None:
            ref = got
        diff = float(np.abs(got - ref).max() / max(1e-30, np.abs(ref).max()))
        lp = mat.last_product()
        print(f"{n}^2 dens {dens:<6} nnz {nnz:>10d} NATIVE={mode}: {lp['kernel']:16s} cold {cold.mean():8.1f} us (min {cold.min():8.1f}) warm {warm:8.1f} us "
              f"{2.0 * nnz / cold.mean() / 1e3:8.1f} GFLOP/s  {lp['algorithmic_bytes'] / cold.mean() / 1e3:7.0f} GB/s of its format  diff {diff:.1e}", flush=True)
    os.environ.pop("SBLAS_B200_SPMV_NATIVE", None)
    del mat, S
